// fly_camera SCRIPT... -- drives tools/gltf_viewer/FlyCamera.hpp (CPU only) with a scripted input sequence and prints the pose after
// every step: "x y z yaw pitch fov vx vy vz".  Steps: place:x,y,z,yaw,pitch,fov | keys:strafe,lift,thrust,look | cursor:x,y |
// scroll:steps | advance:seconds[,repeats].  tests/test_viewer_cpu.py compares the poses with an independent restatement of the
// reference viewer's camera behaviour.
#include "../../tools/gltf_viewer/FlyCamera.hpp"

#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <string>
#include <vector>

int main(int argc, char** argv) {
  auto cam = app::makeFlyCamera(1280, 720);
  app::HeldKeys keys {};
  for (int i = 1; i < argc; ++i) {
    std::string const arg = argv[i];
    auto const colon = arg.find(':');
    std::string const op = arg.substr(0, colon);
    std::vector<double> v;
    std::stringstream ss(colon == std::string::npos ? "" : arg.substr(colon + 1));
    for (std::string tok; std::getline(ss, tok, ',');) v.push_back(std::atof(tok.c_str()));
    if (op == "place" && v.size() == 6) app::place(cam, {float(v[0]), float(v[1]), float(v[2])}, float(v[3]), float(v[4]), float(v[5]));
    else if (op == "keys" && v.size() == 4) keys = app::HeldKeys {int(v[0]), int(v[1]), int(v[2]), v[3] != 0.0};
    else if (op == "cursor" && v.size() == 2) app::cursorTo(cam, keys, v[0], v[1]);
    else if (op == "scroll" && v.size() == 1) app::scrollBy(cam, v[0]);
    else if (op == "advance" && !v.empty()) for (int k = 0; k < (v.size() > 1 ? int(v[1]) : 1); ++k) app::advance(cam, keys, v[0]);
    else { std::fprintf(stderr, "bad step %s\n", arg.c_str()); return 2; }
    std::printf("%.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n", cam.position.x, cam.position.y, cam.position.z, cam.yaw, cam.pitch, cam.fov, cam.velocity.x,
                cam.velocity.y, cam.velocity.z);
  }
  auto const data = app::cameraData(cam);
  std::printf("view");
  for (int c = 0; c < 4; ++c)
    for (int r = 0; r < 4; ++r) std::printf(" %.9g", data.view[c][r]);
  std::printf("\n");
  return 0;
}
