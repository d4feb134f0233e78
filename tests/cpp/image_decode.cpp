// image_decode IN.png OUT.rgba -- decodes a PNG with pbr::image::decode (include/pbr/image/Decode.hpp) and writes the
// raw R8G8B8A8 pixels; prints "width height".  Exit 2 + message on a decode error.  CPU only (no libpbrcu).
#include "pbr/image/Decode.hpp"

#include <cstdio>
#include <fstream>
#include <iterator>

int main(int argc, char** argv) {
  if (argc != 3) {
    std::fprintf(stderr, "usage: image_decode IN.png OUT.rgba\n");
    return 1;
  }
  std::ifstream in(argv[1], std::ios::binary);
  std::vector<std::uint8_t> bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
  try {
    auto const img = pbr::image::decode(bytes);
    std::ofstream out(argv[2], std::ios::binary);
    out.write(reinterpret_cast<char const*>(img.rgba.data()), std::streamsize(img.rgba.size()));
    std::printf("%u %u\n", img.width, img.height);
  } catch (std::exception const& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 2;
  }
  return 0;
}
