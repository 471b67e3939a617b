// Template shapes (S states, I inputs, horizon N) instantiated in this build.  dlib::mpc<S,I,horizon> is a
// compile-time template as well (mpc.h:13-18); each shape is one translation unit (solve_inst.cu compiled with
// -DBMPC_S/-DBMPC_I/-DBMPC_N) so that the build parallelises.  build.py reads this list.
//   (2,2,4)    the reference module's own controller, src/trajectory_point_follower.cpp:366
//   (2,1,30)   dlib's known-answer test, dlib_files/dlib/test/mpc.cpp:295
//   (4,1,N)    BASELINE.json configs, N in {10,30,60,100} (synthetic model)
//   (4,2,10), (3,2,7)  extra shapes for differential tests (I > 1 with S > I, odd sizes)
#pragma once
#define BMPC_SHAPES(X) X(2, 2, 4) X(2, 1, 30) X(4, 1, 10) X(4, 1, 30) X(4, 1, 60) X(4, 1, 100) X(4, 2, 10) X(3, 2, 7)
