// Host-only behaviour tests of the kept pbr:: interfaces (no GPU needed): the parts of the reference
// that are pure CPU code -- Scene/Node traversal order (Scene.cpp:56-61,77-82,97-104), MeshBuilder
// offsets (MeshBuilder.cpp:16-55), makeCameraData (CameraData.hpp:23-32), makeModelPushConstant
// (ModelPushConstant.hpp:33-46).  A heap-backed IAllocator is injected at the seam the reference
// provides (memory/IAllocator.hpp:10-20).  Prints "ok <name>" per case; exit code = failures.
#include "pbr/CameraUniform.hpp"
#include "pbr/MeshBuilder.hpp"
#include "pbr/ModelPushConstant.hpp"
#include "pbr/Scene.hpp"
#include "pbr/ViewBlock.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace {
int failures = 0;
#define EXPECT(cond)                                                        \
  do {                                                                      \
    if (!(cond)) { std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); ++failures; } \
  } while (0)
bool near(float a, float b, float tol = 1e-6f) { return std::fabs(a - b) <= tol * (1.0f + std::fabs(b)); }

class HeapAllocator final : public pbr::IAllocator {
public:
  auto allocateBuffer(std::size_t size, pbr::AllocationInfo) -> pbr::Allocation override {
    return pbr::Allocation {std::calloc(1, size), size, [](void* p) noexcept { std::free(p); }};
  }
  auto allocateImage(vk::Extent2D e, vk::Format f, pbr::AllocationInfo) -> std::pair<pbr::Allocation, std::size_t> override {
    std::size_t const pitch = std::size_t(e.width) * vk::texelBytes(f);
    return {allocateBuffer(pitch * e.height, {}), pitch};
  }
};

void sceneTraversalIsPostOrder() {
  pbr::Scene scene;
  auto& a = scene.addNode(pbr::Node {"a"});
  auto& a1 = a.addChild(pbr::Node {"a1"});
  a1.addChild(pbr::Node {"a1x"});
  a1.addChild(pbr::Node {"a1y"});
  scene.getTopLevelNodes()[0].addChild(pbr::Node {"a2"});
  scene.addNode(pbr::Node {"b"});
  std::string order;
  for (auto const& n : std::as_const(scene).iterateAllNodes()) { order += n.getName().c_str(); order += ' '; }
  EXPECT(order == "a1x a1y a1 a2 a b ");
  std::string mut;
  for (auto& n : scene.iterateAllNodes()) { n.setTransform({.position = {1, 2, 3}}); mut += n.getName().c_str(); }
  EXPECT(mut == "a1xa1ya1a2ab");
  EXPECT(scene.getTopLevelNodes()[1].getTransform().position == (glm::vec3 {1, 2, 3}));
  EXPECT(scene.getTopLevelNodes().size() == 2 && scene.getTopLevelNodes()[0].getChildren().size() == 2);
  std::printf("ok sceneTraversalIsPostOrder\n");
}

void findCameraReturnsFirstInPostOrder() {
  HeapAllocator alloc;
  pbr::Scene scene;
  EXPECT(!scene.findCamera().has_value());
  auto outer = std::make_shared<pbr::CameraUniform>(alloc, pbr::makeCameraData({0, 0, 5}, {0, 0, 0}, 1.0f, 1.0f));
  auto inner = std::make_shared<pbr::CameraUniform>(alloc, pbr::makeCameraData({9, 9, 9}, {0, 0, 0}, 1.0f, 1.0f));
  auto& root = scene.addNode(pbr::Node {"root"});
  root.setCamera(outer);
  scene.getTopLevelNodes()[0].addChild(pbr::Node {"child"}).setCamera(inner);
  auto found = scene.findCamera();
  EXPECT(found.has_value() && found.value() == inner);  // the child is yielded before its parent
  EXPECT(found.value()->get().position == (glm::vec3 {9, 9, 9}));
  inner->set(pbr::makeCameraData({1, 2, 3}, {0, 0, 0}, 1.0f, 1.0f));
  EXPECT(scene.findCamera().value()->get().position == (glm::vec3 {1, 2, 3}));
  std::printf("ok findCameraReturnsFirstInPostOrder\n");
}

void meshBuilderConcatenatesWithRunningOffsets() {
  auto v = [](float x) { return pbr::MeshVertex {{x, 0, 0}, {0, 0, 1}, {1, 0, 0, 1}, {0, 0}}; };
  pbr::MeshBuilder b;
  b.addPrimitive({v(0), v(1), v(2)}, {0, 1, 2}).addPrimitive({v(3), v(4), v(5), v(6)}, {0, 1, 2, 2, 3, 0});
  b.addPrimitive(pbr::MeshBuilder::Primitive {});
  auto const m = b.build();
  EXPECT(m.vertices.size() == 7 && m.indices.size() == 9 && m.primitives.size() == 3);
  EXPECT(m.primitives[0].firstVertex == 0 && m.primitives[0].vertexCount == 3 && m.primitives[0].firstIndex == 0 && m.primitives[0].indexCount == 3);
  EXPECT(m.primitives[1].firstVertex == 3 && m.primitives[1].vertexCount == 4 && m.primitives[1].firstIndex == 3 && m.primitives[1].indexCount == 6);
  EXPECT(m.primitives[2].firstVertex == 7 && m.primitives[2].vertexCount == 0 && m.primitives[2].firstIndex == 9 && m.primitives[2].indexCount == 0);
  EXPECT(m.indices[3] == 0 && m.indices[8] == 0);  // indices stay primitive-local
  EXPECT(m.vertices[3].position.x == 3.0f && m.primitives[1].material == nullptr);
  std::printf("ok meshBuilderConcatenatesWithRunningOffsets\n");
}

void cameraDataMatchesClosedForm() {
  auto const c = pbr::makeCameraData({0, 0, 5}, {0, 0, 0}, 1.5707963f, 16.0f / 9.0f);
  EXPECT(c.position == (glm::vec3 {0, 0, 5}));
  // looking down -Z with up = -Y: s = f x up = (0,0,-1)x(0,-1,0) = (-1,0,0); u = s x f = (0,-1,0)
  EXPECT(near(c.view[0][0], -1) && near(c.view[1][1], -1) && near(c.view[2][2], 1) && near(c.view[3][2], -5));
  glm::vec4 const eye = c.view * glm::vec4 {0, 0, 0, 1};  // world origin -> 5 units in front (-z)
  EXPECT(near(eye.z, -5) && near(eye.x, 0) && near(eye.y, 0));
  float const t = std::tan(1.5707963f / 2);
  EXPECT(near(c.proj[0][0], 1 / (16.0f / 9.0f * t)) && near(c.proj[1][1], 1 / t) && near(c.proj[2][3], -1));
  EXPECT(near(c.proj[2][2], -(1024.0f + 0.01f) / (1024.0f - 0.01f)) && near(c.proj[3][2], -2 * 1024.0f * 0.01f / (1024.0f - 0.01f)));
  glm::vec4 const nearPt = c.proj * glm::vec4 {0, 0, -0.01f, 1}, farPt = c.proj * glm::vec4 {0, 0, -1024.0f, 1};
  EXPECT(near(nearPt.z / nearPt.w, -1.0f, 1e-4f) && near(farPt.z / farPt.w, 1.0f, 1e-4f));  // "NO" clip range
  static_assert(sizeof(pbr::CameraData) == 144 && alignof(pbr::CameraData) == 16);
  std::printf("ok cameraDataMatchesClosedForm\n");
}

void modelPushConstantNormalMatrix() {
  auto const q = glm::angleAxis(0.7f, glm::normalize(glm::vec3 {1, 2, 3}));
  auto const pc = pbr::makeModelPushConstant({1, 2, 3}, q, {2, 3, 4});
  // model * inverse(model) = I
  auto const id = pc.model * glm::inverse(pc.model);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) EXPECT(near(id[i][j], i == j ? 1.0f : 0.0f, 1e-5f));
  // translation column, and a normal transformed by normalModel stays perpendicular to a transformed tangent
  EXPECT(near(pc.model[3].x, 1) && near(pc.model[3].y, 2) && near(pc.model[3].z, 3) && near(pc.model[3].w, 1));
  glm::vec3 const n {0, 0, 1}, tng {1, 0, 0};
  glm::mat3 nm;
  for (int i = 0; i < 3; ++i) nm[i] = glm::vec3 {pc.normalModel[i].x, pc.normalModel[i].y, pc.normalModel[i].z};
  glm::vec3 const n2 = nm * n, t2 = glm::toMat3(pc.model) * tng;
  EXPECT(std::fabs(glm::dot(n2, t2)) < 1e-5f);
  EXPECT(pc.normalModel[0].w == 0.0f && pc.normalModel[2].w == 0.0f);
  // pure rotation: normal matrix == rotation
  auto const r = pbr::makeModelPushConstant({0, 0, 0}, q, {1, 1, 1});
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) EXPECT(near(r.normalModel[i][j], r.model[i][j], 1e-5f));
  std::printf("ok modelPushConstantNormalMatrix\n");
}

void uniformRoundTrips() {
  HeapAllocator alloc;
  pbr::Uniform<pbr::MaterialData> u {alloc, {.color = {0.1f, 0.2f, 0.3f, 1.0f}}};
  EXPECT(u.get().color.y == 0.2f);
  u.set({.color = {4, 5, 6, 7}});
  EXPECT(u.get().color.w == 7.0f && u.getUniformBuffer().size() == sizeof(pbr::MaterialData));
  std::printf("ok uniformRoundTrips\n");
}
}  // namespace

void viewBlockLayout() {
  HeapAllocator alloc;
  pbr::ViewBlock const block {alloc, {.width = 20, .height = 12}, 5, true};
  EXPECT(block.getViewCount() == 5 && block.isExtended() && block.pitch8() == 160 && block.pitch4() == 80);
  EXPECT(block.viewStride8() == 160 * 12 && block.viewStride4() == 80 * 12 && block.frameStride() == block.viewStride8());
  auto const g0 = block.raw(), g3 = block.raw(3);
  EXPECT(g0.width == 20 && g0.height == 12 && g0.pitch8 == 160 && g0.pitch4 == 80);
  EXPECT(static_cast<char*>(g3.positions) - static_cast<char*>(g0.positions) == std::ptrdiff_t(3 * block.viewStride8()));
  EXPECT(static_cast<char*>(g3.emissive) - static_cast<char*>(g0.emissive) == std::ptrdiff_t(3 * block.viewStride8()));
  EXPECT(reinterpret_cast<char*>(g3.depth) - reinterpret_cast<char*>(g0.depth) == std::ptrdiff_t(3 * block.viewStride4()));
  EXPECT(static_cast<char*>(g3.orm) - static_cast<char*>(g0.orm) == std::ptrdiff_t(3 * block.viewStride4()));
  EXPECT(static_cast<char*>(block.getFrame(4).data) - static_cast<char*>(block.getFrames()) == std::ptrdiff_t(4 * block.frameStride()) && block.getFrame(4).pitch == 160);
  pbr::ViewBlock const plain {alloc, {.width = 4, .height = 4}, 2, false};
  EXPECT(plain.raw(1).orm == nullptr && plain.raw(1).emissive == nullptr && plain.raw(1).albedo != nullptr);
  bool threw = false;
  try { pbr::ViewBlock bad {alloc, {.width = 4, .height = 4}, 0, false}; } catch (std::runtime_error const&) { threw = true; }
  EXPECT(threw);
  std::printf("ok viewBlockLayout\n");
}

int main() {
  viewBlockLayout();
  sceneTraversalIsPostOrder();
  findCameraReturnsFirstInPostOrder();
  meshBuilderConcatenatesWithRunningOffsets();
  cameraDataMatchesClosedForm();
  modelPushConstantNormalMatrix();
  uniformRoundTrips();
  std::printf("%s (%d failures)\n", failures ? "FAILED" : "all host facade tests passed", failures);
  return failures;
}
