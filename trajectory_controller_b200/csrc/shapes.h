// Template shapes (S states, I inputs, horizon N) x structural patterns instantiated in this build.
// dlib::mpc<S,I,horizon> is a compile-time template as well (mpc.h:13-18); each (shape, pattern) is one translation
// unit (solve_inst.cu compiled with -DBMPC_S/-DBMPC_I/-DBMPC_N/-DBMPC_PAT) so that the build parallelises.
// list_shapes.py reads this list for the Makefile.
//   (2,2,4)    the reference module's own controller, src/trajectory_point_follower.cpp:366
//   (2,1,30)   dlib's known-answer test, dlib_files/dlib/test/mpc.cpp:295
//   (4,1,N)    BASELINE.json configs, N in {10,30,60,100} (synthetic model)
//   (4,2,10), (3,2,7)  extra shapes for differential tests (I > 1 with S > I, odd sizes)
// Patterns (patterns.h): `dense` accepts every model of the shape and also carries the constructor kernel;
// the named ones drop structurally-zero / exactly-one coefficients for the models this project ships with.
#pragma once
#ifdef BMPC_QUICK   /* developer A/B builds: only the benchmark shape (make QUICK=1 OUT=...) */
#define BMPC_SHAPES(X) X(4, 1, 30, dense) X(4, 1, 30, bicycle4)
#else
#define BMPC_SHAPES(X)                                                                                        \
    X(2, 2, 4, dense) X(2, 1, 30, dense) X(4, 1, 10, dense) X(4, 1, 30, dense) X(4, 1, 60, dense)            \
    X(4, 1, 100, dense) X(4, 2, 10, dense) X(3, 2, 7, dense)                                                  \
    X(2, 2, 4, tobi) X(2, 1, 30, kat) X(4, 1, 10, bicycle4) X(4, 1, 30, bicycle4) X(4, 1, 60, bicycle4)      \
    X(4, 1, 100, bicycle4)
#endif
