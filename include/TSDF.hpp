// TSDF.hpp — header-only C++ shim with the reference's `class TSDF` surface
// (reference include/tsdf.hpp:22-98) over the C ABI of tsdf_b200.h.
//
// Un-commenting the reference's own call sites compiles against this header unchanged:
//   src/Object.cpp:67   tsdf = new TSDF(mnHeight, mnWidth, mnId, base2world, origin);
//   src/Object.cpp:76   delete(tsdf);
//   src/Object.cpp:164  tsdf->Integrate(depth, cam2world_vec);
//   include/Object.hpp:217  TSDF* tsdf;
//
// Behaviour kept: 200^3 grid, 4 mm voxels, trunc = 5 voxels, TUM3 intrinsics
// (include/tsdf.hpp:63-67,96); public host mirrors voxel_grid_TSDF / voxel_grid_weight that
// are refreshed only at destruction (src/tsdf.cu:101-104); destructor writes tsdf<id>.ply
// and tsdf<id>.bin into the working directory (src/tsdf.cu:107-132); a CUDA failure is
// fatal — the reference prints and exit()s (src/tsdf.cu:405-420), this shim throws
// std::runtime_error instead of killing the process.
// Behaviour fixed: device and host memory are released (the reference frees nothing), the
// kernel launch is error-checked, a singular base pose is reported (src/tsdf.cu:74 ignores it).
#ifndef TSDF_B200_TSDF_HPP
#define TSDF_B200_TSDF_HPP

#include <stdexcept>
#include <string>
#include <vector>

#include "tsdf_b200.h"

class TSDF {
 public:
  /// @param h depth image height  @param w depth image width  @param id object id (file names)
  /// @param base2world_vec row-major 4x4 pose of the base frame (first observing keyframe)
  /// @param origin grid origin in the base frame (src/Object.cpp:37-49)
  TSDF(int h, int w, int id, std::vector<float> base2world_vec, std::vector<float> origin, int device = 0)
      : voxel_grid_TSDF(nullptr), voxel_grid_weight(nullptr), vol_(nullptr), mnId(id) {
    if (base2world_vec.size() != 16 || origin.size() != 3)
      throw std::invalid_argument("TSDF: base2world needs 16 floats and origin 3");
    tsdf_config cfg;
    tsdf_config_default(&cfg);
    cfg.im_h = h;
    cfg.im_w = w;
    cfg.id = id;
    cfg.device = device;
    for (int i = 0; i < 16; ++i) cfg.base2world[i] = base2world_vec[i];
    for (int i = 0; i < 3; ++i) cfg.origin[i] = origin[i];
    check(tsdf_create(&cfg, &vol_), "tsdf_create");
    n_ = (size_t)cfg.dim_x * cfg.dim_y * cfg.dim_z;
    voxel_grid_TSDF = new float[n_];
    voxel_grid_weight = new float[n_];
    for (size_t i = 0; i < n_; ++i) voxel_grid_TSDF[i] = 1.0f;   // src/tsdf.cu:79-80
    for (size_t i = 0; i < n_; ++i) voxel_grid_weight[i] = 0.0f;  // src/tsdf.cu:81
  }

  ~TSDF() {
    if (vol_) {
      // src/tsdf.cu:98-133: download, surface points -> tsdf<id>.ply, raw grid -> tsdf<id>.bin
      tsdf_download(vol_, voxel_grid_TSDF, voxel_grid_weight, nullptr, 0);
      const std::string base = "tsdf" + std::to_string(mnId);
      tsdf_save_ply(vol_, (base + ".ply").c_str(), TSDF_PRED_NONZERO, 1.2f, 0.9f);
      tsdf_save_bin(vol_, (base + ".bin").c_str());
      tsdf_destroy(vol_);
    }
    delete[] voxel_grid_TSDF;
    delete[] voxel_grid_weight;
  }

  TSDF(const TSDF&) = delete;
  TSDF& operator=(const TSDF&) = delete;

  /// Integrates one depth image (fp32 metres, h*w, row-major; borrowed for the call) taken at
  /// camera-to-world pose cam2world_vec (16 floats, row-major).  src/tsdf.cu:135-168.
  void Integrate(float* depth_im, std::vector<float> cam2world_vec) {
    if (cam2world_vec.size() != 16) throw std::invalid_argument("TSDF::Integrate: pose needs 16 floats");
    // the reference's upload is blocking (src/tsdf.cu:161-162), so the caller may reuse
    // depth_im immediately; keep that contract
    check(tsdf_integrate(vol_, depth_im, nullptr, nullptr, cam2world_vec.data(), 1.0f), "tsdf_integrate");
  }

  /// Extension: refresh the public mirrors without destroying the object.
  void Download() { check(tsdf_download(vol_, voxel_grid_TSDF, voxel_grid_weight, nullptr, 0), "tsdf_download"); }

  /// Extension: surface points (base frame), the data the reference's viewers were meant to
  /// draw ("to be replaced with TSDF points", src/ObjectDrawer.cpp:35).
  std::vector<float> SurfacePoints(float weight_thresh = 0.9f) {
    int64_t n = 0;
    check(tsdf_extract_points(vol_, TSDF_PRED_NONZERO, 1.2f, weight_thresh, nullptr, nullptr, nullptr, 0, &n), "count");
    std::vector<float> xyz((size_t)n * 3);
    if (n) check(tsdf_extract_points(vol_, TSDF_PRED_NONZERO, 1.2f, weight_thresh, xyz.data(), nullptr, nullptr, n, &n), "extract");
    return xyz;
  }

  tsdf_vol* handle() { return vol_; }

  /// pointer to a TSDF voxel grid (host mirror; stale until Download()/destruction, as in the reference)
  float* voxel_grid_TSDF;
  /// pointer to a TSDF voxel grid weights
  float* voxel_grid_weight;

 private:
  static void check(int rc, const char* what) {
    if (rc != TSDF_OK) throw std::runtime_error(std::string("TSDF: ") + what + ": " + tsdf_last_error());
  }
  tsdf_vol* vol_;
  size_t n_ = 0;
  int mnId;
};

#endif  // TSDF_B200_TSDF_HPP
