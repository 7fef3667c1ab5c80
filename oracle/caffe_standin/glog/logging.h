// TEST INFRASTRUCTURE (oracle). Stand-in for <glog/logging.h> so the reference's bundled Caffe source
// 3rdparty/caffe/src/caffe/util/im2col.cpp compiles alone: its headers only name the CHECK macros inside inline
// helpers the oracle never calls.
#pragma once
#include <iostream>
#define CHECK(x) if (false) std::cerr
#define CHECK_GT(a, b) if (false) std::cerr
#define CHECK_GE(a, b) if (false) std::cerr
#define CHECK_LT(a, b) if (false) std::cerr
#define CHECK_LE(a, b) if (false) std::cerr
#define CHECK_EQ(a, b) if (false) std::cerr
#define CHECK_NE(a, b) if (false) std::cerr
#define LOG(x) if (false) std::cerr
#define DCHECK(x) if (false) std::cerr
#define DCHECK_GT(a, b) if (false) std::cerr
#define DCHECK_GE(a, b) if (false) std::cerr
#define DCHECK_LT(a, b) if (false) std::cerr
#define DCHECK_LE(a, b) if (false) std::cerr
#define DCHECK_EQ(a, b) if (false) std::cerr
