// Exercises the reference-signature overloads of the TSDFfusion shim — Integrate(cv::Mat imRGB,
// cv::Mat imD) (reference include/TSDFfusion.hpp:49) and the 3-argument form with the pose —
// against the raw-pointer overload, on a tiny synthetic frame.  Built with the mock cv::Mat of
// tests/mock_opencv; prints "ok" or exits non-zero.  exit 2 = library error (e.g. no device).
#include <cstdio>
#include <vector>

#include "TSDFfusion.hpp"

int main() {
  const int H = 48, W = 64;
  std::vector<float> depth((size_t)H * W, 1.0f);
  std::vector<uint8_t> rgb((size_t)H * W * 3);
  for (size_t i = 0; i < rgb.size(); ++i) rgb[i] = (uint8_t)(i * 7);
  for (int r = 0; r < H; ++r) depth[(size_t)r * W + (r % W)] = 0.0f;  // some holes
  float pose[16] = {1, 0, 0, 0.01f, 0, 1, 0, -0.02f, 0, 0, 1, 0, 0, 0, 0, 1};
  const float K[9] = {60.f, 0, 32.f, 0, 60.f, 24.f, 0, 0, 1};
  try {
    std::vector<double> bnds = {-0.6, 0.6, -0.5, 0.5, 0.5, 1.3};
    TSDFfusion a(bnds, 0.05, H, W), b(bnds, 0.05, H, W), c(bnds, 0.05, H, W);
    a.SetIntrinsics(K); b.SetIntrinsics(K); c.SetIntrinsics(K);
    a.Integrate(rgb.data(), depth.data(), pose, 1.0f);                    // raw pointers
    cv::Mat imRGB(H, W, CV_8UC3, rgb.data()), imD(H, W, CV_32FC1, depth.data()), T(4, 4, CV_32FC1, pose);
    b.Integrate(imRGB, imD, T);                                           // cv::Mat + pose
    c.SetPose(pose);
    c.Integrate(imRGB, imD);                                              // the reference's 2-argument signature
    std::vector<float> ta, ca, tb, cb, tc, cc;
    a.GetVolume(ta, ca); b.GetVolume(tb, cb); c.GetVolume(tc, cc);
    size_t touched = 0;
    for (size_t i = 0; i < ta.size(); ++i) touched += ta[i] != 1.0f || ca[i] != 0.0f;
    if (ta != tb || ca != cb || ta != tc || ca != cc || touched == 0) { fprintf(stderr, "overloads disagree\n"); return 1; }
    bool threw = false;
    try { cv::Mat bad(H, W, CV_8UC3, rgb.data()); b.Integrate(imRGB, bad, T); } catch (const std::runtime_error&) { threw = true; }
    if (!threw) { fprintf(stderr, "type check missing\n"); return 1; }
  } catch (const std::exception& e) {
    fprintf(stderr, "cvmat_overloads: %s\n", e.what());
    return 2;
  }
  printf("ok\n");
  return 0;
}
