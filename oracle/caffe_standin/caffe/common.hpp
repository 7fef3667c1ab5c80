// TEST INFRASTRUCTURE (oracle). Stand-in for caffe/common.hpp (the real one needs boost, gflags, glog, protobuf): the few
// names math_functions.hpp's declarations use. Shadows the real header because this directory comes first on the include
// path; every other caffe/... header is the reference's own.
#pragma once
#include <cstring>
#include <string>
#include <vector>
#include "glog/logging.h"
namespace caffe
{
    using std::string;
    using std::vector;
}
