// ref_tinyobj.cpp — TEST INFRASTRUCTURE ONLY.
// The reference never defines TINYOBJLOADER_IMPLEMENTATION in any TU, so it does
// not link as checked in.  This two-line TU supplies the implementation of the vendored header,
// included from /root/reference/RaytracingEngine (nothing is copied into this repository).
#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"
