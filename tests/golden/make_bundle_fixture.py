"""Generates tests/golden/bundle_head.txt: the header and the first 4 cameras of the reference's
saved fr3_office run (result/rgbd/bundle.txt), the input format of tsdf_read_bundle_poses.
Run in the build container only (reads /root/reference)."""
import sys
src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/result/rgbd/bundle.txt"
n = 4
with open(src) as f:
    f.readline()
    body = [f.readline() for _ in range(5 * n)]
with open(__file__.replace("make_bundle_fixture.py", "bundle_head.txt"), "w") as o:
    o.write(f"{n} 0\n")
    o.writelines(body)
