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
    // vlcal::VisualCameraCalibration over one NIDCost (visual_camera_calibration.cpp:49-84)
    pcp_params p;
    pcp_default_params(&p);
    p.image_width = p.cull_width = 320;
    p.image_height = p.cull_height = 240;
    pcp::Colorizer col(proj, p, 0);
    std::vector<double> img(320 * 240);
    for (size_t i = 0; i < img.size(); i++) img[i] = (double)((i % 320) / 2 % 160) / 160.0;
    for (size_t i = 0; i < kept.size(); i++) kept[i].intensity = (float)(i % 97) / 97.0f;
    std::vector<std::shared_ptr<vlcal::NIDCost>> costs{std::make_shared<vlcal::NIDCost>(col, img.data(), 320, 240, kept)};
    vlcal::VisualCameraCalibration calib(costs);
    const pcp::Isometry3d T = calib.calibrate(pcp::Isometry3d::Identity());
    const pcp_calib_summary& cs = calib.summary();
    if (!(cs.final_cost <= cs.initial_cost) || !(T.m[15] == 1.0)) { std::printf("calibrate: %f -> %f\n", cs.initial_cost, cs.final_cost); return 14; }
    // one keyframe with a mask through the colouring calls: the per-keyframe mask cloud and the compacted final cloud
    col.setCloud(pts);
    std::vector<uint8_t> bgr(320 * 240 * 3), msk(320 * 240);
    for (size_t i = 0; i < msk.size(); i++) { bgr[3 * i] = (uint8_t)(i % 251); bgr[3 * i + 1] = (uint8_t)(i % 241 + 1); bgr[3 * i + 2] = 7; msk[i] = (i % 5 == 0) ? 255 : (uint8_t)(i % 200); }
    const pcp::Image im{bgr.data(), 320, 240, 3, 320 * 3}, mk{msk.data(), 320, 240, 1, 320};
    pcp_frame_out fo{};
    pcp::Colorizer::keepPixels(fo);
    col.addFrame(pcp::Pose{0, 0, 0, 1, 0, 0, 0}, im, &mk, nullptr, &fo);
    const auto seg = col.segmentCloud();
    const auto coloured = col.coloredCloud();
    if (seg.empty() || seg.size() != (size_t)fo.n_colored || coloured.size() != seg.size()) { std::printf("mask cloud %zu, coloured %zu, frame %lld\n", seg.size(), coloured.size(), (long long)fo.n_colored); return 15; }
    for (const auto& q : seg)
      if (q.a != 255 || (q.segmentMask == 255 && !(q.r == 255 && q.g == 0 && q.b == 0))) { std::printf("mask cloud: bad record\n"); return 16; }
    std::printf("facade ok: kept %zu of %zu, calibrate %d outer / %d evaluations, mask cloud %zu points\n", kept.size(), pts.size(), cs.outer_iterations,
                cs.cost_evaluations, seg.size());
    return kept.empty() ? 11 : 0;
  } catch (const std::runtime_error& e) {
    std::printf("facade error: %s\n", e.what());
    return 3;
  }
}
