// obj_dump_serial.cpp - TEST INFRASTRUCTURE: `obj_dump_serial MODEL.obj PREFIX` dumps the scene a model loads to through
// the line-by-line iostream loader (obj_serial_loader.hpp), in the format of `raytrace --dump-scene`.
// tests/test_host_mirror.py compares the two dumps bit for bit.  Built on demand by the tests with g++; no CUDA.
#include <cstdio>

#include "../../ray-tracing-project_b200/host/scene_dump.hpp"
#include "obj_serial_loader.hpp"

int main(int argc, char **argv) {
  if (argc != 3) {
    fprintf(stderr, "usage: obj_dump_serial MODEL.obj PREFIX\n");
    return 2;
  }
  return scene_dump::dump_scene(argv[1], argv[2], 300, false,
                                [](Tucano::Mesh &m, std::vector<Tucano::Material::Mtl> &mt, const std::string &f) {
                                  Tucano::MeshImporter::loadObjFileSerial(m, mt, f);
                                },
                                "serial");
}
