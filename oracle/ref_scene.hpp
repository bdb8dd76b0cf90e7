// oracle/ref_scene.hpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// The reference's translation unit (found through -I$(REF), nothing copied) and the scene handle the two test
// libraries share: libsrt_ref.so (the reference + restated loops; links NOTHING of the product) and
// libsrt_ref_dropin.so (the product's drop-in headers compiled against the reference's own types; links libsrt_b200).
#pragma once
#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <unistd.h>

#define main reference_main_never_called
#include "simple_raytracer.cpp"   // the reference TU, found through -I$(REF)
#undef main

#include "srt_flatten.hpp"        // product header (duck-typed on the reference's member names)

namespace {
struct RefScene {
    ObjectManager om;
    srt::FlatScene flat;
    bool flat_valid = false;
    // leaf Node* -> first index in the flattened (leaf-order) triangle array
    std::unordered_map<const Node*, int> leaf_base;
    std::vector<std::string> names;                 // objTriangles iteration order
    std::string last_error;
};

}  // namespace
