/* Test-infrastructure shim: classifier.ino expects a header with `float data[]` (the clip under test, not in the
 * reference tree).  The oracle calls the stage functions directly, so a silent second is enough. */
#pragma once
static float data[16000] = {0.0f};
