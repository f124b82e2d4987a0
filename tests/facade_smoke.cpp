// Compile-and-link check of the C++ facade (include/pcp/pcp.hpp) against libpcp_b200.so; run by tests/test_capi_load.py.
// Without a GPU it must fail loudly (no CPU fallback); with one it culls a tiny cloud.
#include <cstdio>
#include <pcp/pcp.hpp>

int main() {
  auto proj = camera::create_camera("pinhole", {376.4, 376.4, 158.8, 122.8}, {0.003, 0.066, -0.0002, -0.0007, 0.0});
  try {
    camera::create_camera("fisheye", {1, 1, 0, 0}, {});
    return 10;
  } catch (const std::runtime_error&) {}
  try {  // PointCloudProcessor: reference constructor signature; a missing PCD throws the reference's message
    PointCloudProcessor proc("/nonexistent/scans.pcd", "/nonexistent/odom.txt", "/nonexistent/", "", "/tmp/", false, false, false);
    proc.process();
    return 12;
  } catch (const std::runtime_error& e) {
    if (std::string(e.what()).find("Couldn't read point cloud file") == std::string::npos) { std::printf("unexpected: %s\n", e.what()); return 13; }
  }
  std::vector<pcp::PointXYZI> pts;
  for (int i = 0; i < 1000; i++) pts.push_back({0.001f * (i % 50) - 0.025f, 0.001f * (i / 50) - 0.01f, 1.0f + 0.01f * (i % 7), 1.0f, 0.5f, 0, 0, 0});
  try {
    vlcal::ViewCulling vc(proj, 320, 240, vlcal::ViewCullingParams());
    auto kept = vc.cull(pts, pcp::Isometry3d::Identity());
    std::printf("facade ok: kept %zu of %zu\n", kept.size(), pts.size());
    return kept.empty() ? 11 : 0;
  } catch (const std::runtime_error& e) {
    std::printf("facade error: %s\n", e.what());
    return 3;
  }
}
