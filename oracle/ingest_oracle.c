/* ingest_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/Makefile): CPU restatement of the two firmware-side
 * conversions that feed classify().  Only tests/ may call it; the product never does.
 *   sj_oracle_adc12            sync/sync.cpp:236-241  int16_t pcmSample = (adc_val - 2048) * 16; buffer[] = pcmSample / 32768.0
 *   sj_oracle_upsample_linear  microphone/src/microphone.cpp:55-70 (upsampleLinear), restated line by line
 * Parity: the firmware sources need the Pico / Particle SDKs and cannot be compiled here, so these are restatements
 * ("parity unpinned" for the target's FMA contraction; here multiply and add are separate, as on x86-64). */
#include <math.h>
#include <stdint.h>

void sj_oracle_adc12(const uint16_t *adc, long n, float *out)
{
    for (long i = 0; i < n; i++) {
        int16_t pcmSample = (int16_t)(((int)adc[i] - 2048) * 16);
        out[i] = (float)(pcmSample / 32768.0);
    }
}

void sj_oracle_upsample_linear(const float *inBuffer, int oldSize, float *outBuffer, int newSize)
{
    for (int i = 0; i < newSize; i++) {
        float oldIndex = i * ((float)(oldSize - 1) / (float)(newSize - 1));
        int indexFloor = (int)floor(oldIndex);
        int indexCeil = (indexFloor == (oldSize - 1)) ? (oldSize - 1) : (indexFloor + 1);
        float frac = oldIndex - indexFloor;
        float valFloor = inBuffer[indexFloor];
        float valCeil = inBuffer[indexCeil];
        outBuffer[i] = valFloor + (valCeil - valFloor) * frac;
    }
}
