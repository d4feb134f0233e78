// gltf_viewer -- the reference's viewer executable (src/gltf_viewer/Main.cpp:15-28: `gltf_viewer <asset> [-vulkan-validation]`),
// headless: renders a fixed number of frames of the asset's first scene into device images and writes the last
// frame (PNG / PFM) and the per-frame times (JSON) instead of presenting to a window.
//   gltf_viewer <asset.glb|asset.gltf> [-vulkan-validation] [--size WxH] [--frames N] [--dt SECONDS]
//               [--keys WASDQE] [--look DX,DY] [--scroll Y] [--position X,Y,Z] [--yaw R] [--pitch R] [--fov R]
//               [--no-spin] [--ggx] [--out frame.png] [--hdr frame.pfm] [--times times.json] [--dump DIR]
//               [--views N [--view-first K] [--view-total M] [--orbit RADIUS,HEIGHT]]   (CUDA addition: a block of orbit views of the
//                scene through PbrRenderSystem::renderViews -- BASELINE configs[4] -- instead of the frame loop; --frames = timed repeats)
// No arguments: exits 0 without doing anything, like the reference.
#include "App.hpp"

#include <cstdlib>
#include <span>
#include <string_view>

namespace {
auto parseFloats(std::string const& text, int n, char sep) -> std::vector<double> {
  std::vector<double> out;
  std::size_t pos = 0;
  while (pos <= text.size()) {
    auto const end = text.find(sep, pos);
    out.push_back(std::stod(text.substr(pos, end == std::string::npos ? std::string::npos : end - pos)));
    if (end == std::string::npos) break;
    pos = end + 1;
  }
  if (int(out.size()) != n) throw std::runtime_error("expected " + std::to_string(n) + " values in '" + text + "'");
  return out;
}
}  // namespace

auto main(int const argc, char const* const* const argv) -> int {
  std::span const args(argv + 1, std::size_t(argc - 1));
  if (args.empty()) return 0;
  try {
    app::Options opt;
    opt.asset = args.front();
    for (std::size_t i = 1; i < args.size(); ++i) {
      std::string_view const a = args[i];
      auto const next = [&]() -> std::string {
        if (i + 1 >= args.size()) throw std::runtime_error("missing value after " + std::string(a));
        return args[++i];
      };
      if (a == "-vulkan-validation") continue;  // accepted for command-line compatibility; there is no Vulkan instance
      else if (a == "--size") { auto const v = parseFloats(next(), 2, 'x'); opt.extent = {std::uint32_t(v[0]), std::uint32_t(v[1])}; }
      else if (a == "--frames") opt.frames = std::uint32_t(std::stoul(next()));
      else if (a == "--dt") opt.fixedDeltaTime = std::stod(next());
      else if (a == "--keys") {
        for (char const c : next()) {
          switch (c) {
            case 'W': case 'w': opt.input.thrust = 1; break;
            case 'A': case 'a': opt.input.strafe = -1; break;
            case 'S': case 's': if (opt.input.thrust == 0) opt.input.thrust = -1; break;
            case 'D': case 'd': if (opt.input.strafe == 0) opt.input.strafe = 1; break;
            case 'Q': case 'q': if (opt.input.lift == 0) opt.input.lift = -1; break;
            case 'E': case 'e': opt.input.lift = 1; break;
            default: throw std::runtime_error("unknown key in --keys");
          }
        }
      }
      else if (a == "--look") { auto const v = parseFloats(next(), 2, ','); opt.lookX = v[0]; opt.lookY = v[1]; opt.input.look = true; }
      else if (a == "--scroll") opt.scroll = std::stod(next());
      else if (a == "--position") { auto const v = parseFloats(next(), 3, ','); opt.position = glm::vec3 {float(v[0]), float(v[1]), float(v[2])}; }
      else if (a == "--yaw") opt.yaw = std::stof(next());
      else if (a == "--pitch") opt.pitch = std::stof(next());
      else if (a == "--fov") opt.fov = std::stof(next());
      else if (a == "--no-spin") opt.spin = false;
      else if (a == "--ggx") opt.ggx = true;
      else if (a == "--out") opt.outPng = next();
      else if (a == "--hdr") opt.outPfm = next();
      else if (a == "--times") opt.outTimes = next();
      else if (a == "--dump") opt.dumpDir = next();
      else if (a == "--views") opt.views = std::uint32_t(std::stoul(next()));
      else if (a == "--view-first") opt.viewFirst = std::uint32_t(std::stoul(next()));
      else if (a == "--view-total") opt.viewTotal = std::uint32_t(std::stoul(next()));
      else if (a == "--orbit") { auto const v = parseFloats(next(), 2, ','); opt.orbitRadius = float(v[0]); opt.orbitHeight = float(v[1]); }
      else throw std::runtime_error("unknown option " + std::string(a));
    }
    if (opt.extent.width == 0 || opt.extent.height == 0 || opt.frames == 0) throw std::runtime_error("--size and --frames must be positive");
    bool const viewsMode = opt.views > 0;
    app::App app(std::move(opt));
    if (viewsMode) {  // a block of orbit views through the batched passes (--frames = timed repeats, default 60: pass a small number)
      app.runViews();
      return 0;
    }
    app.run();
    double sum = 0.0;
    for (auto const t : app.frameTimes()) sum += t;
    std::printf("Frame time %.3f ms\n", sum / double(app.frameTimes().size()));
  } catch (std::exception const& e) {
    std::fprintf(stderr, "gltf_viewer: %s\n", e.what());
    return 1;
  }
  return 0;
}
