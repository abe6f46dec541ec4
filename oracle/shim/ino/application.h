/* Test-infrastructure shim: the minimum of the Particle/Arduino environment that
 * donut-classifier/classifier.ino needs to compile on a host: PI and a Serial object.  Serial does not print;
 * it records what the sketch reports (numbers in order, and whether the "We have a Scrub Jay" line was printed)
 * so that tests can read the verdict of the unmodified setup().  No classifier logic here. */
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifndef PI
#define PI 3.1415926535897932384626433832795
#endif
extern "C" {
extern double sj_shim_serial_numbers[4096];
extern int sj_shim_serial_count;
extern int sj_shim_serial_scrub; /* 1 after "We have a Scrub Jay! :)", 0 after "We have no Scrub Jay! :(" */
}
struct SjShimSerial {
    void note(const char *s)
    {
        if (strstr(s, "We have a Scrub Jay")) sj_shim_serial_scrub = 1;
        else if (strstr(s, "We have no Scrub Jay")) sj_shim_serial_scrub = 0;
    }
    void num(double v) { if (sj_shim_serial_count < 4096) sj_shim_serial_numbers[sj_shim_serial_count++] = v; }
    void print(const char *s) { note(s); }
    void println(const char *s) { note(s); }
    void print(int v) { num(v); }
    void println(int v) { num(v); }
    void print(float v) { num(v); }
    void println(float v) { num(v); }
    void print(double v) { num(v); }
    void println(double v) { num(v); }
    void begin(int) {}
};
static SjShimSerial Serial;
