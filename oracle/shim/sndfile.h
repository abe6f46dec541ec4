/* shim: donut-classifier/classifier.c includes <sndfile.h> (line 9) but calls no sf_* function; it
 * only relies on the header's transitive includes for int16_t / off_t.  Test infrastructure. */
#ifndef SJ_SHIM_SNDFILE_H
#define SJ_SHIM_SNDFILE_H
#include <stdint.h>
#include <sys/types.h>
#endif
