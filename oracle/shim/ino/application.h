/* Test-infrastructure shim: the minimum of the Particle/Arduino environment that
 * donut-classifier/classifier.ino needs to compile on a host (a Serial sink and PI).  No logic. */
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifndef PI
#define PI 3.1415926535897932384626433832795
#endif
struct SjShimSerial {
    template <typename T> void print(const T &) {}
    template <typename T> void println(const T &) {}
    void begin(int) {}
};
static SjShimSerial Serial;
