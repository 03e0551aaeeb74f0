"""The C++ host side of the drop-in boundary: include/TSDF.hpp + include/TSDFfusion.hpp (the
reference's class surfaces, reference include/tsdf.hpp:22-98 and include/TSDFfusion.hpp:25-65)
compiled with g++ into examples/replay_rgbd — the keyframe loop of the reference's
examples/label_instance_rgbd.cpp:79-110 — and, on the GPU box, its output files compared
with the oracle byte for byte."""
import os
import struct
import subprocess

import numpy as np
import pytest

from conftest import ROOT, bits
from oracle.oracle import SEM_TSDFCU, SEM_FUSIONPY

FACTOR = 5000.0
BNDS = np.array([[-0.7, 0.7], [-0.6, 0.6], [0.7, 2.0]])
FUSION_VS = 0.02


@pytest.fixture(scope="module")
def replay_bin(pkg, tmp_path_factory):
    out = tmp_path_factory.mktemp("cpp") / "replay_rgbd"
    libdir = os.path.dirname(pkg.capi.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "examples", "replay_rgbd.cpp"), "-o", str(out),
           "-L", libdir, "-ltsdf_b200", f"-Wl,-rpath,{libdir}"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return str(out)


def make_sequence(synth, path, n=3):
    scene = synth.Scene().sphere((0, 0, 1.3), 0.3).plane((0, 0, 1), 1.8)
    rng = np.random.default_rng(17)
    poses = [synth.pose(synth.rot_xyz(0.01 * i, -0.015 * i, 0.005 * i), (0.012 * i, -0.004 * i, 0)) for i in range(n)]
    frames = []
    with open(path, "wb") as f:
        f.write(struct.pack("<4i", 0x31515354, n, *synth.TUM3_HW))
        f.write(synth.TUM3_K.astype(np.float32).tobytes())
        f.write(struct.pack("<f", FACTOR))
        f.write(BNDS.astype(np.float32).tobytes())
        f.write(struct.pack("<f", FUSION_VS))
        for T in poses:
            d, rgb = synth.render(scene, synth.TUM3_K, synth.TUM3_HW, T, rng)
            u16 = np.clip(np.round(d * FACTOR), 0, 65535).astype(np.uint16)
            f.write(np.ascontiguousarray(T, np.float32).tobytes())
            f.write(u16.tobytes())
            f.write(np.ascontiguousarray(rgb, np.uint8).tobytes())
            # what the host loop hands to Integrate (examples/label_instance_rgbd.cpp:89-100)
            sub = np.zeros(synth.TUM3_HW, np.float32)
            sub[::4, ::3] = u16[::4, ::3].astype(np.float32) * (np.float32(1.0) / np.float32(FACTOR))
            frames.append((sub, rgb, np.ascontiguousarray(T, np.float32)))
    return frames


def test_cpp_host_builds_and_fails_loudly_without_inputs(replay_bin, tmp_path):
    r = subprocess.run([replay_bin], capture_output=True, text=True)
    assert r.returncode == 1 and "usage" in r.stderr
    r = subprocess.run([replay_bin, str(tmp_path / "missing.bin"), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1


def test_cpp_host_has_no_cpu_fallback(replay_bin, synth, pkg, tmp_path):
    """Without a CUDA device the program must stop with the library's error, not compute."""
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    make_sequence(synth, tmp_path / "seq.bin", n=1)
    r = subprocess.run([replay_bin, str(tmp_path / "seq.bin"), str(tmp_path), "tsdf"], capture_output=True, text=True)
    assert r.returncode == 2 and "replay_rgbd:" in r.stderr
    assert not (tmp_path / "tsdf7.bin").exists()


@pytest.mark.gpu
@pytest.mark.parametrize("raw", [False, True])
def test_cpp_replay_matches_oracle(replay_bin, synth, oracle, tmp_path, raw):
    frames = make_sequence(synth, tmp_path / "seq.bin")
    args = [replay_bin, str(tmp_path / "seq.bin"), str(tmp_path), "tsdf" if raw else "both"] + (["--raw"] if raw else [])
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    # class TSDF path: 200^3 @ 4 mm, base = first Twc, origin by src/Object.cpp:37-49
    dims, vs = (200, 200, 200), 0.004
    base = frames[0][2]
    origin = oracle.instance_origin(frames[0][0], None, synth.TUM3_K)
    p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), origin, synth.TUM3_HW, synth.TUM3_K, 6.0, SEM_TSDFCU)
    t, w, _ = oracle.new_volume(p)
    for d, _, T in frames:
        oracle.integrate(p, oracle.compose(base, T), d, None, 1.0, t, w, None)
    assert w.max() > 0
    assert np.array_equal(np.fromfile(tmp_path / "weights.f32", np.float32), w)
    assert oracle.save_bin(str(tmp_path / "o.bin"), p, t) and oracle.save_ply(str(tmp_path / "o.ply"), p, t, w)
    assert (tmp_path / "o.bin").read_bytes() == (tmp_path / "tsdf7.bin").read_bytes()
    assert (tmp_path / "o.ply").read_bytes() == (tmp_path / "tsdf7.ply").read_bytes()
    if raw:
        return
    # class TSDFfusion path: dims = ceil(extent / vs), colour + obs_weight
    b32 = BNDS.astype(np.float32).astype(np.float64)
    d3 = tuple(int(v) for v in np.ceil((b32[:, 1] - b32[:, 0]) / np.float64(np.float32(FUSION_VS))))
    vsf = float(np.float32(FUSION_VS))
    p2 = oracle.params(d3, vsf, np.float32(vsf) * np.float32(5), b32[:, 0], synth.TUM3_HW, synth.TUM3_K, np.inf,
                       SEM_FUSIONPY)
    t2, w2, c2 = oracle.new_volume(p2, True)
    for d, rgb, T in frames:
        oracle.integrate(p2, T.reshape(16), d, oracle.pack_color(rgb), 1.0, t2, w2, c2)
    assert np.array_equal(bits(np.fromfile(tmp_path / "fusion_tsdf.f32", np.float32)), bits(t2))
    assert np.array_equal(np.fromfile(tmp_path / "fusion_color.f32", np.float32), c2)
    assert oracle.save_ply(str(tmp_path / "o2.ply"), p2, t2, w2)
    assert (tmp_path / "o2.ply").read_bytes() == (tmp_path / "fusion_points.ply").read_bytes()
    # SaveMesh (src/TSDFfusion.py.in:48-53): marching cubes + meshwrite, against the oracle's bytes
    mv, mf, mn, mc = oracle.mesh(p2, t2, c2)
    assert len(mf) > 100 and oracle.save_mesh_ply(str(tmp_path / "o2_mesh.ply"), mv, mf, mn, mc)
    assert (tmp_path / "o2_mesh.ply").read_bytes() == (tmp_path / "fusion_mesh.ply").read_bytes()


@pytest.mark.gpu
def test_cpp_replay_on_a_device_list_writes_the_same_files(replay_bin, synth, pkg, tmp_path):
    """`TSDFfusion(bnds, vs, h, w, devices)`: the C++ host drives N z-slabs in ONE process (no Python,
    no torchrun in the loop) and every output file — volume, colour, point cloud, mesh — equals the
    single-device run's, byte for byte.  [0,0,0] on a 1-GPU box, across devices where there are more."""
    make_sequence(synth, tmp_path / "seq.bin")
    lists = ["0,0,0"] + (["0,1", "1,0,1"] if pkg.device_count() >= 2 else [])
    ref = tmp_path / "single"
    ref.mkdir()
    r = subprocess.run([replay_bin, str(tmp_path / "seq.bin"), str(ref), "fusion"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for k, devs in enumerate(lists):
        out = tmp_path / f"slabs{k}"
        out.mkdir()
        r = subprocess.run([replay_bin, str(tmp_path / "seq.bin"), str(out), "fusion", "--devices", devs],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        for name in ("fusion_tsdf.f32", "fusion_color.f32", "fusion_points.ply", "fusion_mesh.ply"):
            assert (out / name).read_bytes() == (ref / name).read_bytes(), f"{devs}: {name} differs"
        assert (out / "fusion_mesh.ply").stat().st_size > 10000


@pytest.fixture(scope="module")
def cvmat_bin(pkg, tmp_path_factory):
    """tests/cpp/cvmat_overloads.cpp against the mock <opencv2/core.hpp> (OpenCV headers are absent
    from the container): the -DTSDF_B200_WITH_OPENCV overloads of include/TSDFfusion.hpp must compile."""
    out = tmp_path_factory.mktemp("cvm") / "cvmat_overloads"
    libdir = os.path.dirname(pkg.capi.LIB_PATH)
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-DTSDF_B200_WITH_OPENCV",
           "-I", os.path.join(ROOT, "tests", "mock_opencv"), "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "cvmat_overloads.cpp"), "-o", str(out),
           "-L", libdir, "-ltsdf_b200", f"-Wl,-rpath,{libdir}"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return str(out)


def test_cvmat_overloads_compile_and_fail_loudly_without_device(cvmat_bin, pkg):
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    r = subprocess.run([cvmat_bin], capture_output=True, text=True)
    assert r.returncode == 2 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_cvmat_overloads_agree_with_raw_pointers(cvmat_bin):
    """Integrate(cv::Mat imRGB, cv::Mat imD) — the reference's signature — and the pose-taking form
    give the volume of the raw-pointer call; wrong Mat types throw std::runtime_error."""
    r = subprocess.run([cvmat_bin], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "ok", r.stderr
