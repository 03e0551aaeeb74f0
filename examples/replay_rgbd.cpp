// replay_rgbd.cpp — C++ host program on the drop-in shims (include/TSDF.hpp, TSDFfusion.hpp).
//
// It is the keyframe loop of the reference's off-line replay
// (reference examples/label_instance_rgbd.cpp:79-110) reduced to the TSDF hot path: for every
// keyframe read depth + colour + Twc, subsample the depth 3x4 (:89-97), scale by
// 1/DepthMapFactor (:99-100) and integrate it — into one `TSDF` object set up the way
// Object::Object does (base pose = first keyframe's Twc, origin = min of the back-projected
// depth, reference src/Object.cpp:23-49) and/or into a `TSDFfusion` colour volume.
// ORB-SLAM2, Mask-RCNN and the viewers are out of scope, so the frames come from one flat
// sequence file instead of a TUM directory:
//
//   int32  magic 'TSQ1', n_frames, H, W
//   float  K[9], DepthMapFactor, fusion vol_bnds[6], fusion voxel_size
//   per frame: float Twc[16]; uint16 depth[H*W]; uint8 rgb[H*W*3]
//
// usage: replay_rgbd <sequence.bin> <out_dir> [tsdf|fusion|both] [--raw | --devices 0,1,...]
//   tsdf   -> <out_dir>/tsdf<id>.ply, tsdf<id>.bin (written by ~TSDF, src/tsdf.cu:107-132),
//             weights.f32 (the public mirror voxel_grid_weight)
//   fusion -> <out_dir>/fusion_tsdf.f32, fusion_color.f32, fusion_points.ply, fusion_mesh.ply
//             (SaveMesh of the plugin, src/TSDFfusion.py.in:48-53)
//   --raw  -> hand the u16 image to tsdf_integrate_raw_async (subsampling + scaling on the GPU)
//   --devices a,b,.. -> the TSDFfusion volume as z-slabs over that device list, one process, no Python
//             (a device may repeat); the output files must not change
// exit codes: 0 ok, 1 usage / file error, 2 library error (e.g. no CUDA device: no CPU fallback)
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "TSDF.hpp"
#include "TSDFfusion.hpp"

namespace {

struct Sequence {
  int32_t n = 0, H = 0, W = 0;
  float K[9], factor = 5000.0f, bnds[6], fusion_vs = 0.02f;
  std::vector<std::vector<float>> poses;
  std::vector<std::vector<uint16_t>> depth;
  std::vector<std::vector<uint8_t>> rgb;
};

bool load(const char* path, Sequence& s) {
  FILE* f = fopen(path, "rb");
  if (!f) return false;
  int32_t hdr[4];
  bool ok = fread(hdr, 4, 4, f) == 4 && hdr[0] == 0x31515354 /* 'TSQ1' */;
  if (ok) {
    s.n = hdr[1], s.H = hdr[2], s.W = hdr[3];
    ok = fread(s.K, 4, 9, f) == 9 && fread(&s.factor, 4, 1, f) == 1 && fread(s.bnds, 4, 6, f) == 6 &&
         fread(&s.fusion_vs, 4, 1, f) == 1;
  }
  const size_t px = (size_t)s.H * s.W;
  for (int i = 0; ok && i < s.n; ++i) {
    s.poses.emplace_back(16);
    s.depth.emplace_back(px);
    s.rgb.emplace_back(px * 3);
    ok = fread(s.poses[i].data(), 4, 16, f) == 16 && fread(s.depth[i].data(), 2, px, f) == px &&
         fread(s.rgb[i].data(), 1, px * 3, f) == px * 3;
  }
  fclose(f);
  return ok;
}

// examples/label_instance_rgbd.cpp:89-100: keep col%3==0 && row%4==0, then scale to metres.
void preprocess(const Sequence& s, const std::vector<uint16_t>& raw, std::vector<float>& out) {
  const float scale = 1.0f / s.factor;
  out.assign((size_t)s.H * s.W, 0.0f);
  for (int col = 0; col < s.W; col += 3)
    for (int row = 0; row < s.H; row += 4) out[(size_t)row * s.W + col] = (float)raw[(size_t)row * s.W + col] * scale;
}

// src/Object.cpp:37-49: origin = componentwise min of the back-projected valid pixels, start 1000.
std::vector<float> object_origin(const Sequence& s, const std::vector<float>& depth) {
  std::vector<float> o(3, 1000.0f);
  const float invfx = 1.0f / s.K[0], invfy = 1.0f / s.K[4], cx = s.K[2], cy = s.K[5];
  for (int r = 0; r < s.H; ++r)
    for (int c = 0; c < s.W; ++c) {
      const float z = depth[(size_t)r * s.W + c];
      if (z > 0) {
        const float x = (c - cx) * z * invfx, y = (r - cy) * z * invfy;
        if (x < o[0]) o[0] = x;
        if (y < o[1]) o[1] = y;
        if (z < o[2]) o[2] = z;
      }
    }
  return o;
}

void dump(const std::string& path, const float* p, size_t n) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f || fwrite(p, 4, n, f) != n) throw std::runtime_error("cannot write " + path);
  fclose(f);
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <sequence.bin> <out_dir> [tsdf|fusion|both] [--raw]\n", argv[0]);
    return 1;
  }
  const std::string mode = argc > 3 ? argv[3] : "both";
  const bool raw = argc > 4 && !strcmp(argv[4], "--raw");
  std::vector<int> devices(1, 0);
  if (argc > 5 && !strcmp(argv[4], "--devices")) {
    devices.clear();
    for (const char* p = argv[5]; *p;) {
      devices.push_back(atoi(p));
      while (*p && *p != ',') ++p;
      if (*p == ',') ++p;
    }
  }
  Sequence s;
  if (!load(argv[1], s) || s.n < 1) {
    fprintf(stderr, "cannot read sequence %s\n", argv[1]);
    return 1;
  }
  if (chdir(argv[2]) != 0) {  // ~TSDF writes into the working directory, as the reference does
    fprintf(stderr, "cannot enter %s\n", argv[2]);
    return 1;
  }
  try {
    std::vector<float> depth;
    if (mode == "tsdf" || mode == "both") {
      const int id = 7;
      preprocess(s, s.depth[0], depth);
      std::unique_ptr<TSDF> tsdf(new TSDF(s.H, s.W, id, s.poses[0], object_origin(s, depth)));  // src/Object.cpp:67
      for (int i = 0; i < s.n; ++i) {
        printf("Processing frame %d:\n", i);
        if (raw) {
          int rc = tsdf_integrate_raw_async(tsdf->handle(), s.depth[i].data(), TSDF_DEPTH_U16, 1.0f / s.factor, 3, 4,
                                            nullptr, nullptr, s.poses[i].data(), 1.0f);
          if (rc == TSDF_OK) rc = tsdf_sync(tsdf->handle());
          if (rc != TSDF_OK) throw std::runtime_error(tsdf_last_error());
        } else {
          preprocess(s, s.depth[i], depth);
          tsdf->Integrate(depth.data(), s.poses[i]);  // src/Object.cpp:164
        }
      }
      tsdf->Download();
      dump("weights.f32", tsdf->voxel_grid_weight, (size_t)200 * 200 * 200);
      tsdf.reset();  // src/Object.cpp:76 -> tsdf7.ply, tsdf7.bin
    }
    if (mode == "fusion" || mode == "both") {
      std::vector<double> bnds(s.bnds, s.bnds + 6);
      TSDFfusion fusion(bnds, (double)s.fusion_vs, s.H, s.W, devices);
      fusion.SetIntrinsics(s.K);
      for (int i = 0; i < s.n; ++i) {
        preprocess(s, s.depth[i], depth);
        fusion.Integrate(s.rgb[i].data(), depth.data(), s.poses[i].data(), 1.0f);  // src/TSDFfusion.py.in:43
      }
      std::vector<float> t, c;
      fusion.GetVolume(t, c);
      dump("fusion_tsdf.f32", t.data(), t.size());
      dump("fusion_color.f32", c.data(), c.size());
      fusion.SavePointCloud("fusion_points.ply");
      fusion.SaveMesh("fusion_mesh.ply");  // src/TSDFfusion.py.in:48-53
    }
  } catch (const std::exception& e) {
    fprintf(stderr, "replay_rgbd: %s\n", e.what());
    return 2;
  }
  return 0;
}
