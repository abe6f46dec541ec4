/* Test-infrastructure shim: classifier.ino expects a header with `float data[]`, the clip under test (not in the
 * reference tree).  Here it is a writable one-second buffer that tests fill before calling the sketch's setup(). */
#pragma once
extern "C" {
float data[16000] = {0.0f};
double sj_shim_serial_numbers[4096];
int sj_shim_serial_count = 0;
int sj_shim_serial_scrub = -1;
}
