// CameraController -- the viewer's fly camera, headless.  Mirrors src/gltf_viewer/CameraController.hpp:21-173:
// the same state (position, velocity, pitch, yaw, sensitivity, speed, aspect ratio, fov), constants and update
// arithmetic; what it read from the vkfw window (held keys, right mouse button, framebuffer size) is a plain
// Input record that the frame loop fills from the command line instead of from window events.
#pragma once
#include "pbr/CameraData.hpp"

#include <algorithm>
#include <cmath>
#include <cstddef>

namespace app {
struct Input {          // the window state CameraController polls (vkfw::Window::getKey / getMouseButton)
  bool a {}, d {}, e {}, q {}, w {}, s {};
  bool rightButton {};
};

class CameraController {
public:
  static constexpr auto DEFAULT_POSITION = glm::vec3(0.0f);
  static constexpr auto DEFAULT_PITCH = 0.0f;
  static constexpr auto DEFAULT_YAW = -glm::half_pi<float>();
  static constexpr auto DEFAULT_SENSITIVITY = 0.01f;
  static constexpr auto DEFAULT_SPEED = 0.02f;
  static constexpr auto DEFAULT_FOV = glm::half_pi<float>();
  static constexpr auto MIN_PITCH = -glm::half_pi<float>() + 0.01f;
  static constexpr auto MAX_PITCH = -MIN_PITCH;
  static constexpr auto MIN_FOV = 0.1f;
  static constexpr auto MAX_FOV = glm::half_pi<float>();
  static constexpr auto ACCELERATION_FACTOR = 10.0f;
  static constexpr auto FOV_INCREMENT = 0.1f;

private:
  Input _input {};
  glm::vec2 _lastMousePos {};
  glm::vec3 _position = DEFAULT_POSITION;
  glm::vec3 _velocity {};
  float _pitch = DEFAULT_PITCH;
  float _yaw = DEFAULT_YAW;
  float _sensitivity = DEFAULT_SENSITIVITY;
  float _speed = DEFAULT_SPEED;
  float _aspectRatio;
  float _fov = DEFAULT_FOV;

public:
  CameraController(std::size_t width, std::size_t height) : _aspectRatio(static_cast<float>(width) / static_cast<float>(height)) {}

  auto setInput(Input input) noexcept -> void { _input = input; }
  // start state (the reference always starts from the DEFAULT_ constants; a headless run has no other way to look around)
  auto setPose(glm::vec3 position, float yaw, float pitch, float fov) noexcept -> void {
    _position = position; _yaw = yaw;
    _pitch = std::clamp(pitch, MIN_PITCH, MAX_PITCH);
    _fov = std::clamp(fov, MIN_FOV, MAX_FOV);
  }

  auto onWindowResize(std::size_t width, std::size_t height) -> void { _aspectRatio = static_cast<float>(width) / static_cast<float>(height); }
  auto onCursorMove(double newX, double newY) -> void {
    glm::vec2 const newMousePos {static_cast<float>(newX), static_cast<float>(newY)};
    glm::vec2 const mouseDelta = newMousePos - _lastMousePos;
    _lastMousePos = newMousePos;
    if (_input.rightButton) {
      _yaw -= mouseDelta.x * _sensitivity;
      _pitch -= mouseDelta.y * _sensitivity;
      _pitch = std::clamp(_pitch, MIN_PITCH, MAX_PITCH);
    }
  }
  auto onScroll(double /*xOffset*/, double yOffset) -> void {
    _fov = std::clamp(_fov - (static_cast<float>(yOffset) * FOV_INCREMENT), MIN_FOV, MAX_FOV);
  }
  auto update(double deltaTime) -> void {
    auto const inputVector = getInputVector();
    if (glm::length(inputVector) != 0.0f) accelerate(glm::normalize(inputVector) * _speed, static_cast<float>(deltaTime) * ACCELERATION_FACTOR);
    _position += getMatrix() * _velocity;
    accelerate({}, static_cast<float>(deltaTime) * ACCELERATION_FACTOR);
  }
  [[nodiscard]] auto getCameraData() const noexcept -> pbr::CameraData {
    return pbr::makeCameraData(_position, _position + getDirection(), _fov, _aspectRatio);
  }
  [[nodiscard]] auto getPosition() const noexcept -> glm::vec3 { return _position; }

private:
  [[nodiscard]] auto getDirection() const noexcept -> glm::vec3 {
    auto const cosPitch = std::cos(_pitch);
    return glm::normalize(glm::vec3(cosPitch * std::cos(_yaw), std::sin(_pitch), cosPitch * std::sin(_yaw)));
  }
  auto accelerate(glm::vec3 acceleration, float lerp) noexcept -> void { _velocity = glm::mix(_velocity, acceleration, lerp); }
  [[nodiscard]] auto getInputVector() const noexcept -> glm::vec3 {
    glm::vec3 vec {};
    if (_input.a) vec.x = -1.0f; else if (_input.d) vec.x = 1.0f;
    if (_input.e) vec.y = 1.0f; else if (_input.q) vec.y = -1.0f;
    if (_input.w) vec.z = 1.0f; else if (_input.s) vec.z = -1.0f;
    return vec;
  }
  [[nodiscard]] auto getMatrix() const noexcept -> glm::mat3x3 {
    auto const forward = getDirection();
    auto const right = glm::cross(forward, {0.0f, 1.0f, 0.0f});
    auto const xyz = [](glm::vec4 v) { return glm::vec3 {v.x, v.y, v.z}; };
    glm::vec3 const upVector = xyz(glm::rotate(glm::mat4 {1.0f}, glm::half_pi<float>(), right) * glm::vec4(forward, 0.0f));
    glm::mat3x3 m;
    m[0] = xyz(glm::rotate(glm::mat4 {1.0f}, glm::half_pi<float>(), upVector) * glm::vec4(forward, 0.0f));
    m[1] = upVector;
    m[2] = forward;
    return m;
  }
};
}  // namespace app
