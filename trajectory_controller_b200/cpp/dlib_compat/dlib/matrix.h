// Drop-in include shim: puts bmpc's fixed-size matrix where the reference module expects dlib's
// (`#include <dlib/matrix.h>`), so that a module written against dlib compiles unchanged when this
// directory replaces `dlib_files` on the include path (reference CMakeLists.txt:12).
#pragma once
#include "bmpc/matrix.h"
namespace dlib {
using bmpc::matrix;
using bmpc::trans;
}  // namespace dlib
