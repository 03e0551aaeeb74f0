"""Generates tests/golden/bundle_head.txt: the header and the first 4 cameras of the reference's
saved fr3_office run (result/rgbd/bundle.txt), the input format of tsdf_read_bundle_poses, and
tests/golden/associations_head.txt: the first 4 lines of result/rgbd/associations.txt
(tsdf_read_associations).
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

with open(src.replace("bundle.txt", "associations.txt")) as f, \
        open(__file__.replace("make_bundle_fixture.py", "associations_head.txt"), "w") as o:
    o.writelines([f.readline() for _ in range(n)])
