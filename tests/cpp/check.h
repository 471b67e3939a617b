// Tiny assertion helper for the C++ boundary tests (no test framework in this image).
#pragma once
#include <cstdio>
#include <cstdlib>
static int g_failed = 0, g_checked = 0;
#define CHECK(cond)                                                                          \
    do {                                                                                     \
        ++g_checked;                                                                         \
        if (!(cond)) { ++g_failed; std::fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); } \
    } while (0)
#define CHECK_MSG(cond, ...)                                                                 \
    do {                                                                                     \
        ++g_checked;                                                                         \
        if (!(cond)) { ++g_failed; std::fprintf(stderr, "FAIL %s:%d: %s  ", __FILE__, __LINE__, #cond); std::fprintf(stderr, __VA_ARGS__); std::fprintf(stderr, "\n"); } \
    } while (0)
static int finish(const char* name)
{
    std::printf("%s: %d checks, %d failed\n", name, g_checked, g_failed);
    return g_failed ? 1 : 0;
}
