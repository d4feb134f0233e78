// FlyCamera -- the headless viewer's free-flight camera: a plain state record advanced by free functions.
// BEHAVIOUR follows the reference's viewer so that a scripted camera path reproduces its frames (SURVEY.md 8(f) next-4:
// src/gltf_viewer/CameraController.hpp -- start pose, limits and rates are the reference's constants, listed in Limits
// below; yaw / pitch from cursor motion while the look button is held, field of view from the scroll wheel, velocity eased
// towards the held direction and back to rest, motion in the camera's own right / up / forward frame).  The code is this
// repository's own: no window, no event callbacks -- the frame loop feeds it the scripted input of the command line --
// and the camera frame comes from two cross products instead of matrix rotations.  tests/test_viewer_cpu.py pins the poses
// after scripted input sequences against an independent restatement.
#pragma once
#include "pbr/CameraData.hpp"

#include <algorithm>
#include <cmath>

namespace app {

// the reference viewer's constants (CameraController.hpp:25-44), as one table
struct Limits {
  static constexpr float startYaw = -1.57079632679489661923f, startPitch = 0.0f, startFov = 1.57079632679489661923f;
  static constexpr float lookRate = 0.01f;                                   // radians per cursor unit
  static constexpr float cruiseSpeed = 0.02f;                                // scene units per update at full velocity
  static constexpr float maxPitch = 1.57079632679489661923f - 0.01f;         // never look straight up or down
  static constexpr float minFov = 0.1f, maxFov = 1.57079632679489661923f, fovPerScrollStep = 0.1f;
  static constexpr float easing = 10.0f;                                     // velocity blend per second
};

// what is held down during a frame: one signed axis per direction (a key pair), and the look button
struct HeldKeys {
  int strafe {};  // -1 = A (left), +1 = D (right)
  int lift {};    // -1 = Q, +1 = E
  int thrust {};  // -1 = S (back), +1 = W (forward)
  bool look {};   // right mouse button: cursor motion turns the camera
};

struct FlyCamera {
  glm::vec3 position {0.0f};
  glm::vec3 velocity {0.0f};  // in the camera's own frame: x right, y up, z forward
  float yaw {Limits::startYaw}, pitch {Limits::startPitch}, fov {Limits::startFov};
  float aspect {1.0f};
  glm::vec2 cursor {0.0f};    // last cursor position seen
};

inline auto makeFlyCamera(std::size_t width, std::size_t height) -> FlyCamera {
  FlyCamera cam;
  cam.aspect = static_cast<float>(width) / static_cast<float>(height);
  return cam;
}

inline auto place(FlyCamera& cam, glm::vec3 position, float yaw, float pitch, float fov) -> void {
  cam.position = position;
  cam.yaw = yaw;
  cam.pitch = std::clamp(pitch, -Limits::maxPitch, Limits::maxPitch);
  cam.fov = std::clamp(fov, Limits::minFov, Limits::maxFov);
}

// unit view direction of (yaw, pitch): yaw turns about +Y starting from +X, pitch lifts towards +Y
inline auto viewDirection(FlyCamera const& cam) -> glm::vec3 {
  float const c = std::cos(cam.pitch);
  return glm::normalize(glm::vec3 {c * std::cos(cam.yaw), std::sin(cam.pitch), c * std::sin(cam.yaw)});
}

// the cursor moved to (x, y): turns the camera when the look button is held, always remembers the position
inline auto cursorTo(FlyCamera& cam, HeldKeys const& keys, double x, double y) -> void {
  glm::vec2 const now {static_cast<float>(x), static_cast<float>(y)};
  glm::vec2 const moved = now - cam.cursor;
  cam.cursor = now;
  if (!keys.look) return;
  cam.yaw -= moved.x * Limits::lookRate;
  cam.pitch = std::clamp(cam.pitch - moved.y * Limits::lookRate, -Limits::maxPitch, Limits::maxPitch);
}

inline auto scrollBy(FlyCamera& cam, double steps) -> void {
  cam.fov = std::clamp(cam.fov - static_cast<float>(steps) * Limits::fovPerScrollStep, Limits::minFov, Limits::maxFov);
}

// one update of `seconds`: ease the velocity towards the held direction, move, ease it back towards rest
inline auto advance(FlyCamera& cam, HeldKeys const& keys, double seconds) -> void {
  float const blend = static_cast<float>(seconds) * Limits::easing;
  glm::vec3 const wish {static_cast<float>(keys.strafe), static_cast<float>(keys.lift), static_cast<float>(keys.thrust)};
  if (glm::length(wish) != 0.0f) cam.velocity = glm::mix(cam.velocity, glm::normalize(wish) * Limits::cruiseSpeed, blend);
  // the camera's frame in world space: forward, `up` a quarter turn from forward about the horizontal axis forward x (+Y), and the
  // sideways axis a quarter turn from forward about `up`
  glm::vec3 const forward = viewDirection(cam);
  glm::vec3 const up = glm::cross(glm::normalize(glm::cross(forward, glm::vec3 {0.0f, 1.0f, 0.0f})), forward);
  glm::vec3 const side = glm::cross(up, forward);
  cam.position += side * cam.velocity.x + up * cam.velocity.y + forward * cam.velocity.z;
  cam.velocity = glm::mix(cam.velocity, glm::vec3 {0.0f}, blend);
}

inline auto cameraData(FlyCamera const& cam) -> pbr::CameraData {
  return pbr::makeCameraData(cam.position, cam.position + viewDirection(cam), cam.fov, cam.aspect);
}
}  // namespace app
