// Everything in front of the reference's TEST_CASE("PhyloHMM") in the file integration/compat/build_catch.sh
// assembles: the reference's Catch header (with its main), Eigen, and the class names of test/test.cpp:10-23 mapped
// onto the adapter. The TEST_CASE itself, and the closing brace of `namespace test` behind it, are the reference's
// own lines, read from /root/reference/test/test.cpp at build time and never stored in this repository.
#define CATCH_CONFIG_MAIN

#include <map>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include <Eigen/Dense>
#include "catch.hpp"
#include "linearham_compat.hpp"

namespace test {

using linearham_compat::GermlineType;
using linearham_compat::PhyloHMM;
using linearham_compat::PhyloHMMPtr;
