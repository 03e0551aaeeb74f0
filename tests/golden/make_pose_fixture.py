"""Generates tests/golden/fr3_office_Twc.npy from the reference's saved fr3_office run.

Run in the build container only (reads /root/reference/result/rgbd/bundle.txt).  Format of
bundle.txt (written by src/ORB_SLAM2/System.cc:884-917, parsed by
examples/label_instance_rgbd.cpp:300-330): header "<n_cams> <n_pts>", then per camera one
intrinsics line, three rotation rows and one translation row of Tcw.  We store Twc = Tcw^-1
(what Engine::Run would pass: KF->GetPoseInverse(), src/Engine.cpp:222) as float32 [n,16].
"""
import sys
import numpy as np

src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/result/rgbd/bundle.txt"
with open(src) as f:
    n_cam, _ = [int(v) for v in f.readline().split()]
    out = np.zeros((n_cam, 16), np.float32)
    for i in range(n_cam):
        f.readline()
        R = np.array([[float(v) for v in f.readline().split()] for _ in range(3)])
        t = np.array([float(v) for v in f.readline().split()])
        Twc = np.eye(4)
        Twc[:3, :3] = R.T
        Twc[:3, 3] = -R.T @ t
        out[i] = Twc.reshape(16).astype(np.float32)
np.save(__file__.replace("make_pose_fixture.py", "fr3_office_Twc.npy"), out)
print(out.shape, out[1].reshape(4, 4))
