// Drop-in include shim for `#include <dlib/control.h>` (reference include/trajectory_point_follower.h:19):
// dlib::mpc<S,I,horizon> becomes the GPU-backed facade of bmpc/mpc.h (same method set as
// dlib_files/dlib/control/mpc.h:13-240).  Nothing else of dlib/control.h (lspi, approximate_linear_models) is
// used by the module, so nothing else is provided.
#pragma once
#include "dlib/matrix.h"
#include "bmpc/mpc.h"
namespace dlib {
using bmpc::mpc;
}  // namespace dlib
