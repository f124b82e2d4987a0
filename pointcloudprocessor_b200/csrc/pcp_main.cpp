// pcp_colorize — command line of the reference's executable (src/main.cpp:7-71) over the drop-in PointCloudProcessor class.
// Same long and short options (boost::program_options is not available here: parsed by hand), same exit codes:
// 0 done, 1 help, -1 missing required arguments, -2 exception.  Extra options lift what the reference hard-codes.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>

#include <pcp/pcp.hpp>

static const char* kHelp =
    "Allowed options:\n"
    "  -h [ --help ]                         Produce help message\n"
    "  -p [ --point_cloud_path ] arg         Path to the point cloud data file\n"
    "  -o [ --odometry_path ] arg            Path to odometry data file\n"
    "  -i [ --images_folder ] arg            Path to directory containing images\n"
    "  -m [ --mask_image_folder ] arg        Path to directory for segmented images\n"
    "  -t [ --output_path ] arg (=.)         Path to save processed output\n"
    "  --enableMLS arg (=0)                  Enable MLS smoothing (not available in this build)\n"
    "  --enableNIDOptimize arg (=0)          Enable NID-based camera pose optimization\n"
    "  --enableInitialGuessManual arg (=0)   Enable manual pickup point based camera pose optimization (not available)\n"
    "  --cull_size WxH                       image_size handed to ViewCulling (reference: 4096x3000)\n"
    "  --intrinsics fx,fy,cx,cy              pinhole intrinsics (reference values by default)\n"
    "  --distortion k1,k2,p1,p2,k3           plumb_bob coefficients (reference values by default)\n"
    "  --device N                            CUDA device\n";

static bool as_bool(const std::string& v) { return v == "1" || v == "true" || v == "on" || v == "yes"; }

int main(int argc, char** argv) {
  try {
    static const std::map<std::string, std::string> shorts = {{"-p", "point_cloud_path"}, {"-o", "odometry_path"}, {"-i", "images_folder"},
                                                              {"-m", "mask_image_folder"}, {"-t", "output_path"}, {"-e", "enableMLS"}};
    std::map<std::string, std::string> opt;
    opt["mask_image_folder"] = "";
    opt["output_path"] = ".";
    for (int i = 1; i < argc; i++) {
      std::string a = argv[i], key, val;
      if (a == "-h" || a == "--help") { std::cout << kHelp << "\n"; return 1; }
      if (a.rfind("--", 0) == 0) {
        const size_t eq = a.find('=');
        key = a.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
        if (eq != std::string::npos) val = a.substr(eq + 1);
        else if (i + 1 < argc) val = argv[++i];
      } else if (shorts.count(a)) {
        key = shorts.at(a);
        if (i + 1 < argc) val = argv[++i];
      } else {
        throw std::runtime_error("unrecognised option '" + a + "'");
      }
      opt[key] = val;
    }
    if (opt.count("point_cloud_path") && opt.count("odometry_path") && opt.count("images_folder")) {
      PointCloudProcessor processor(opt["point_cloud_path"], opt["odometry_path"], opt["images_folder"], opt["mask_image_folder"], opt["output_path"],
                                    as_bool(opt["enableMLS"]), as_bool(opt["enableNIDOptimize"]), as_bool(opt["enableInitialGuessManual"]));
      if (opt.count("cull_size")) {
        int w = 0, h = 0;
        if (std::sscanf(opt["cull_size"].c_str(), "%dx%d", &w, &h) != 2) throw std::runtime_error("--cull_size expects WxH");
        processor.setCullSize(w, h);
      }
      if (opt.count("intrinsics") || opt.count("distortion")) {
        pcp_camera cam{4818.200388954926, 4819.10345841615, 2032.4178620390019, 1535.1895959282901,
                       {0.003043514741045163, 0.06634739187544138, -0.000217681797407554, -0.0006654964142658197, 0.0}};
        if (opt.count("intrinsics") && std::sscanf(opt["intrinsics"].c_str(), "%lf,%lf,%lf,%lf", &cam.fx, &cam.fy, &cam.cx, &cam.cy) != 4)
          throw std::runtime_error("--intrinsics expects fx,fy,cx,cy");
        if (opt.count("distortion") && std::sscanf(opt["distortion"].c_str(), "%lf,%lf,%lf,%lf,%lf", &cam.dist[0], &cam.dist[1], &cam.dist[2],
                                                   &cam.dist[3], &cam.dist[4]) != 5)
          throw std::runtime_error("--distortion expects k1,k2,p1,p2,k3");
        processor.setCamera(cam);
      }
      if (opt.count("device")) processor.setDevice(std::atoi(opt["device"].c_str()));
      processor.process();
      const pcp_process_stats& s = processor.stats();
      std::cout << "points " << s.n_points << ", keyframes " << s.n_keyframes << " of " << s.n_frames << " frames, coloured points " << s.n_colored_points
                << "; load " << s.seconds_load << " s, cull pass " << s.seconds_cull_pass << " s, colour pass " << s.seconds_color_pass
                << " s, writing " << s.seconds_write << " s" << std::endl;
      std::cout << "Processing completed successfully." << std::endl;
    } else {
      std::cerr << "Error: Missing required arguments." << std::endl;
      std::cerr << kHelp << std::endl;
      return -1;
    }
  } catch (const std::exception& e) {
    std::cerr << "Unhandled Exception reached the top of main: " << e.what() << ", application will now exit" << std::endl;
    return -2;
  }
  return 0;
}
