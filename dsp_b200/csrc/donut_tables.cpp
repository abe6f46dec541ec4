// donut_tables.cpp -- host-side constants of the donut kernel.  The window, U and the dB threshold are
// computed with the reference's own formulas on the host libm so that they carry the same bits as
// the reference's run-time values (classifier.c:484-530 / classifier.cpp:263-305).
#include <math.h>
#include <string.h>

#include "donut_kernel.h"

void sj_donut_default_params(SjDonutParams *p, int variant)
{
    // 0 donut-classifier/classifier.c, 1 sync/lib/classifier.cpp,
    // 2 donut-classifier/revert.c = microphone/src/classifier.c (:713, :164/169), 3 microphone/src/classifier.cpp (:79-80, :123, :448, :535)
    p->gap_tol = 0.05; p->min_dur = 0.15; p->lo_norm = 0.70; p->hi_norm = 0.85; p->mid_thr_db = 45.0;
    if (variant == 0)      { p->thr_mid = 75; p->thr_above = 300; p->thr_below = 100; }
    else if (variant == 1) { p->lo_norm = 0.65; p->hi_norm = 0.80; p->mid_thr_db = 70.0; p->thr_mid = 100; p->thr_above = 200; p->thr_below = 80; }
    else if (variant == 2) { p->gap_tol = 0.15; p->thr_mid = 50; p->thr_above = 200; p->thr_below = 200; }
    else if (variant == 3) { p->thr_mid = 100; p->thr_above = 200; p->thr_below = 150; }
    else                   { p->thr_mid = 75; p->thr_above = 300; p->thr_below = 100; } // 4: classifier.ino:88-89, :460, :549-550, :134
    const double b256[3][2] = {{5000, 7000}, {2500, 5000}, {500, 2500}}, bino[3][2] = {{4500, 7500}, {3500, 4000}, {500, 3000}};
    for (int b = 0; b < 3; ++b)
        for (int k = 0; k < 2; ++k) p->band[b][k] = variant == 4 ? bino[b][k] : b256[b][k];
    p->half[0] = 0.18; p->half[1] = 0.05; p->half[2] = 0.18;
    p->nperseg = variant == 4 ? 64 : 256;
}

namespace {

const double kPi = 3.14159265358979323846;
// scipy.signal.butter(4, [lo, hi], 'band', fs=16000) as hard-coded in classifier.c:342-360,383-401
const double kB[2][9] = {{0.01020948, 0., -0.04083792, 0., 0.06125688, 0., -0.04083792, 0., 0.01020948},
                         {0.1362017, 0., -0.5448068, 0., 0.8172102, 0., -0.5448068, 0., 0.1362017}};
const double kA[2][9] = {{1., -4.56803686, 9.95922498, -13.49912589, 12.43979269, -7.94997696, 3.43760562, -0.92305481, 0.1203896},
                         {1., 2.60935592, 2.32553038, 1.20262614, 1.11690211, 0.76154474, 0.10005124, -0.0129829, 0.02236815}};

inline float tukey_rise(float n) { return 0.5f * (1.0f + cosf((float)kPi * (-1.0f + 2.0f * n / (0.25f * (257.0f - 1.0f))))); }
inline float tukey_fall(float n) { return 0.5f * (1.0f + cosf((float)kPi * (-2.0f / 0.25f + 1.0f + 2.0f * n / (0.25f * (257.0f - 1.0f))))); }
inline double tukey_rise(double n) { return 0.5 * (1 + cos(kPi * (-1 + 2.0 * n / (0.25 * (257 - 1))))); }
inline double tukey_fall(double n) { return 0.5 * (1 + cos(kPi * (-2.0 / 0.25 + 1 + 2.0 * n / (0.25 * (257 - 1))))); }

inline bool db_exceeds(float s, float thr) { return (float)(10 * log10(s / 1e-12)) > thr; }
inline bool db_exceeds(double s, double thr) { return 10 * log10(s / 1e-12) > thr; }

template <typename T, typename U> T smallest_exceeding(T thr)
{
    // positive IEEE values are ordered like their bit patterns; dB(s) is monotone in s
    U lo = 1, hi;
    T top;
    if (sizeof(T) == 4) { hi = (U)0x7f7fffffu; } else { hi = (U)0x7fefffffffffffffull; }
    memcpy(&top, &hi, sizeof top);
    if (!db_exceeds(top, thr)) return top; // never exceeded
    while (lo < hi) {
        U mid = lo + (hi - lo) / 2;
        T v;
        memcpy(&v, &mid, sizeof v);
        if (db_exceeds(v, thr)) hi = mid; else lo = mid + 1;
    }
    T v;
    memcpy(&v, &lo, sizeof v);
    return v;
}

// smallest positive a with pred(a) true, pred monotone (false ... false true ... true); max value if never
template <typename T, typename U, typename P> T smallest_with(P pred)
{
    U lo = 1, hi;
    T top;
    if (sizeof(T) == 4) { hi = (U)0x7f7fffffu; } else { hi = (U)0x7fefffffffffffffull; }
    memcpy(&top, &hi, sizeof top);
    if (!pred(top)) return top;
    while (lo < hi) {
        U mid = lo + (hi - lo) / 2;
        T v;
        memcpy(&v, &mid, sizeof v);
        if (pred(v)) hi = mid; else lo = mid + 1;
    }
    T v;
    memcpy(&v, &lo, sizeof v);
    return v;
}

template <typename T, typename U> void build(SjDonutTables<T> *t, const SjDonutParams &p, int fs)
{
    memset(t, 0, sizeof *t);
    for (int n = 0; n <= 32; n++) t->window[n] = tukey_rise((T)n);
    for (int n = 33; n <= 257 - 32 - 2; n++) t->window[n] = 1;
    for (int n = 257 - 32 - 1; n < DN_WIN; n++) t->window[n] = tukey_fall((T)n);
    T u = 0;
    for (int i = 0; i < DN_WIN; i++) u += t->window[i] * t->window[i];
    t->U = u * (T)fs;
    t->fs = (T)fs;
    for (int k1 = 0; k1 < 32; k1++)
        for (int g = 0; g < 4; g++) {
            double ang = 2 * kPi * (g * k1) / 128.0;
            t->tw128[k1][g][0] = (T)cos(ang);
            t->tw128[k1][g][1] = (T)(-sin(ang));
        }
    for (int k = 0; k <= 128; k++) {
        double ang = 2 * kPi * k / 256.0;
        t->tw256[k][0] = (T)cos(ang);
        t->tw256[k][1] = (T)(-sin(ang));
    }
    {   // the twiddle sequence PlainFFT::Compute generates at run time, in its own float arithmetic:
        // u <- u * c per butterfly offset, c <- half-angle of c per stage (sqrt evaluated in double)
        float c1 = -1.0f, c2 = 0.0f;
        for (int l = 0, l1 = 1; l < 8; ++l, l1 <<= 1) {
            float u1 = 1.0f, u2 = 0.0f;
            for (int m = 0; m < l1; ++m) {
                t->bourke[l1 - 1 + m][0] = (T)u1;
                t->bourke[l1 - 1 + m][1] = (T)u2;
                const float z = (u1 * c1) - (u2 * c2);
                u2 = (u1 * c2) + (u2 * c1);
                u1 = z;
            }
            c2 = (float)sqrt((1.0 - c1) / 2.0);
            c2 = -c2; // forward transform
            c1 = (float)sqrt((1.0 + c1) / 2.0);
        }
    }
    for (int w = 0; w < 2; w++)
        for (int j = 0; j < 9; j++) { t->a[w][j] = (T)kA[w][j]; t->b[w][j] = (T)kB[w][j]; }
    t->s_thr = smallest_exceeding<T, U>((T)p.mid_thr_db);
    {
        const volatile T Uv = t->U, thr = t->s_thr; // volatile: keep the divisions as written (no reciprocal tricks)
        t->a_thr[0] = smallest_with<T, U>([&](T a) { T s = a / Uv; return s >= thr; });
        t->a_thr[1] = smallest_with<T, U>([&](T a) { T s = a / Uv; s = s * (T)2; return s >= thr; });
        t->a_pos = smallest_with<T, U>([&](T a) { T s = a / Uv; return s > (T)0; });
        // sum_intense's index search (classifier.c:596-617 / classifier.cpp:372-393) for the three fixed bands
        const T band[3][2] = {{5000, 7000}, {2500, 5000}, {500, 2500}};
        for (int b = 0; b < 3; ++b) {
            int f0 = 0, f1 = DN_BINS - 1;
            while (f0 < DN_BINS && (T)f0 * t->fs / (T)DN_WIN < band[b][0]) f0++;
            while (f1 >= 0 && (T)f1 * t->fs / (T)DN_WIN > band[b][1]) f1--;
            if (f0 >= DN_BINS) f0 = DN_BINS - 1;
            if (f1 < 0) f1 = 0;
            if (f0 > f1) { const int x = f0; f0 = f1; f1 = x; }
            t->box_f[b][0] = f0;
            t->box_f[b][1] = f1;
        }
    }
    t->lo_norm = (T)p.lo_norm; t->hi_norm = (T)p.hi_norm;
    t->gap_tol = (T)p.gap_tol; t->min_dur = (T)p.min_dur;
    t->thr_mid = (T)p.thr_mid; t->thr_above = (T)p.thr_above; t->thr_below = (T)p.thr_below;
}

} // namespace

void sj_donut_build_tables_f32(SjDonutTables<float> *t, const SjDonutParams &p, int fs) { build<float, uint32_t>(t, p, fs); }
void sj_donut_build_tables_f64(SjDonutTables<double> *t, const SjDonutParams &p, int fs) { build<double, uint64_t>(t, p, fs); }

// ---- stage tables (donut_stages.h)
#include "donut_stages.h"

namespace {

template <typename T> T tukey_cos(T x);
template <> float tukey_cos<float>(float x) { return cosf(x); }
template <> double tukey_cos<double>(double x) { return cos(x); }

// the reference's window code (classifier.cpp:263-297 / classifier.c:484-521 / classifier.ino:286-320) for any
// window_size, alpha = 0.25; the write to window[window_size] (out of bounds in the reference) is not made
template <typename T> void build_stage(SjStageTab<T> *t, int nperseg, int fs)
{
    memset(t, 0, sizeof *t);
    t->nperseg = nperseg;
    t->hop = nperseg - nperseg / 8;
    t->bins = nperseg / 2 + 1;
    const T alpha = (T)0.25, M = (T)(nperseg + 1), pi = (T)kPi;
    const int width = (int)floor((double)(alpha * (M - (T)1) / (T)2));
    for (int n = 0; n <= width && n < nperseg; n++)
        t->window[n] = (T)0.5 * ((T)1 + tukey_cos<T>(pi * ((T)-1 + (T)2 * (T)n / (alpha * (M - (T)1)))));
    for (int n = width + 1; n <= (int)(M - width - 2) && n < nperseg; n++) t->window[n] = 1;
    for (int n = (int)(M - width - 1); n < nperseg; n++)
        t->window[n] = (T)0.5 * ((T)1 + tukey_cos<T>(pi * ((T)-2 / alpha + (T)1 + (T)2 * (T)n / (alpha * (M - (T)1)))));
    T u = 0;
    for (int i = 0; i < nperseg; i++) u += t->window[i] * t->window[i];
    t->U = u * (T)fs;
    if (sizeof(T) == 4) {
        // PlainFFT::Compute's run-time twiddle recurrence (sync/lib/PlainFFT.cpp:48-80), float arithmetic
        float c1 = -1.0f, c2 = 0.0f;
        for (int l1 = 1; l1 < nperseg; l1 <<= 1) {
            float u1 = 1.0f, u2 = 0.0f;
            for (int m = 0; m < l1; ++m) {
                t->tw[l1 - 1 + m][0] = (T)u1;
                t->tw[l1 - 1 + m][1] = (T)u2;
                const float z = (u1 * c1) - (u2 * c2);
                u2 = (u1 * c2) + (u2 * c1);
                u1 = z;
            }
            c2 = -(float)sqrt((1.0 - c1) / 2.0);
            c1 = (float)sqrt((1.0 + c1) / 2.0);
        }
    } else {
        for (int l1 = 1; l1 < nperseg; l1 <<= 1)
            for (int m = 0; m < l1; ++m) {
                const double ang = kPi * m / l1;
                t->tw[l1 - 1 + m][0] = (T)cos(ang);
                t->tw[l1 - 1 + m][1] = (T)(-sin(ang));
            }
    }
}

} // namespace

void sj_stage_build_tab_f32(SjStageTab<float> *t, int nperseg, int fs) { build_stage<float>(t, nperseg, fs); }
void sj_stage_build_tab_f64(SjStageTab<double> *t, int nperseg, int fs) { build_stage<double>(t, nperseg, fs); }

void sj_stage_transition(const double a[9], int steps, double M[64])
{
    // zero-input step: s'[0] = -sum_j a[j] s[j-1], s'[k] = s[k-1]; M = A^steps by binary powering (long double)
    long double A[64], R[64], Tm[64];
    for (int i = 0; i < 64; i++) { A[i] = 0; R[i] = 0; }
    for (int j = 1; j < 9; j++) A[j - 1] = -(long double)a[j];
    for (int k = 1; k < 8; k++) A[k * 8 + k - 1] = 1;
    for (int i = 0; i < 8; i++) R[i * 8 + i] = 1;
    auto mul = [&](const long double *X, const long double *Y, long double *Z) {
        for (int i = 0; i < 8; i++)
            for (int j = 0; j < 8; j++) {
                long double acc = 0;
                for (int k = 0; k < 8; k++) acc += X[i * 8 + k] * Y[k * 8 + j];
                Z[i * 8 + j] = acc;
            }
    };
    for (int e = steps; e > 0; e >>= 1) {
        if (e & 1) { mul(R, A, Tm); memcpy(R, Tm, sizeof R); }
        mul(A, A, Tm);
        memcpy(A, Tm, sizeof A);
    }
    for (int i = 0; i < 64; i++) M[i] = (double)R[i];
}

// ---- generic epilogue constants (donut_stages.h)
namespace {
template <typename T, typename U> void build_epilogue(SjEpilogueParams<T> *e, int variant, int fs, int max_mid)
{
    SjDonutParams p;
    sj_donut_default_params(&p, variant);
    memset(e, 0, sizeof *e);
    e->fs = (T)fs;
    e->nperseg = p.nperseg;
    e->hop = p.nperseg - p.nperseg / 8;
    e->bins = p.nperseg / 2 + 1;
    e->max_mid = max_mid;
    e->s_thr = smallest_exceeding<T, U>((T)p.mid_thr_db);
    e->lo_norm = (T)p.lo_norm; e->hi_norm = (T)p.hi_norm; e->gap_tol = (T)p.gap_tol; e->min_dur = (T)p.min_dur;
    e->thr_mid = (T)p.thr_mid; e->thr_above = (T)p.thr_above; e->thr_below = (T)p.thr_below;
    for (int b = 0; b < 3; ++b) {
        e->half[b] = (T)p.half[b];
        // sum_intense's index search (classifier.c:596-617 / classifier.cpp:372-393) on frequencies[i] = i * fs / nfft
        const int bins = e->bins;
        int f0 = 0, f1 = bins - 1;
        while (f0 < bins && (T)f0 * e->fs / (T)p.nperseg < (T)p.band[b][0]) f0++;
        while (f1 >= 0 && (T)f1 * e->fs / (T)p.nperseg > (T)p.band[b][1]) f1--;
        if (f0 >= bins) f0 = bins - 1;
        if (f1 < 0) f1 = 0;
        if (f0 > f1) { const int x = f0; f0 = f1; f1 = x; }
        e->box_f[b][0] = f0;
        e->box_f[b][1] = f1;
    }
}
} // namespace

bool sj_epilogue_variant_f64(int variant) { return variant == 0 || variant == 2; }
int sj_epilogue_variant_nperseg(int variant) { return variant == 4 ? 64 : 256; }
void sj_epilogue_build_params_f32(SjEpilogueParams<float> *p, int variant, int fs, int max_mid) { build_epilogue<float, uint32_t>(p, variant, fs, max_mid); }
void sj_epilogue_build_params_f64(SjEpilogueParams<double> *p, int variant, int fs, int max_mid) { build_epilogue<double, uint64_t>(p, variant, fs, max_mid); }
