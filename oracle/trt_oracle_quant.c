/*
 * trt_oracle_quant.c -- CPU ORACLE for the terminal quantiser (test infrastructure only; PARITY
 * pinned only by the source text: the quantiser is Rust and stays a hand restatement, see trt_oracle.h).  Restates /root/reference/src/lib.rs:503-532: one row of ANSI text
 * per image row, rows concatenated in order.  f32 arithmetic, glibc powf/sqrtf (= Rust f32::powf /
 * f32::sqrt on Linux), Rust `as u8` / `as usize` cast semantics (truncate, saturate, NaN -> 0),
 * Rust f32::clamp (NaN stays NaN) and f32::min (NaN.min(a) == a).
 */
#include "trt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* lib.rs:521 -- 68 glyphs, darkest to brightest */
static const char RAMP[] = " .`^\",:;Il!i><~+_-?][}{1)(|\\tfjrxnuvczXYUJCLQ0OZmwqpdbkhao*#MW&8%B@$";
_Static_assert(sizeof(RAMP) - 1 == 68, "ramp must hold 68 glyphs");

static inline float rust_clamp(float v, float lo, float hi) {   /* f32::clamp */
    if (v < lo) return lo;
    if (v > hi) return hi;
    return v;                                                    /* NaN falls through */
}
static inline uint32_t rust_as_u8(float v) {                     /* `as u8` */
    if (!(v > 0.0f)) return 0u;                                  /* NaN, negatives, zero */
    if (v >= 255.0f) return 255u;
    return (uint32_t)v;
}
static inline float rust_min(float a, float b) {                 /* f32::min */
    if (isnan(a)) return b;
    if (isnan(b)) return a;
    return a < b ? a : b;
}
static inline size_t rust_as_usize(float v) {
    if (!(v > 0.0f)) return 0;
    if (v >= 18446744073709551616.0f) return (size_t)-1;
    return (size_t)v;
}

static size_t put_u8(uint8_t *dst, uint32_t v) {                 /* `{}` of a u8 */
    size_t n = 0;
    if (v >= 100u) { dst[n++] = (uint8_t)('0' + v / 100u); v %= 100u; dst[n++] = (uint8_t)('0' + v / 10u); v %= 10u; }
    else if (v >= 10u) { dst[n++] = (uint8_t)('0' + v / 10u); v %= 10u; }
    dst[n++] = (uint8_t)('0' + v);
    return n;
}

/* channel values before the integer cast; mode 0: lib.rs:515-523, mode 1: lib.rs:510-512 */
static void pre_cast(const float *px, int mode, float *r, float *g, float *b, float *ci) {
    if (mode == 1) {
        *r = rust_clamp(sqrtf(px[0]) * 255.0f, 0.0f, 255.0f);
        *g = rust_clamp(sqrtf(px[1]) * 255.0f, 0.0f, 255.0f);
        *b = rust_clamp(sqrtf(px[2]) * 255.0f, 0.0f, 255.0f);
        *ci = 0.0f;
    } else {
        const float gamma = 0.3f;
        *r = rust_clamp(powf(px[0], gamma) * 255.0f, 0.0f, 255.0f);
        *g = rust_clamp(powf(px[1], gamma) * 255.0f, 0.0f, 255.0f);
        *b = rust_clamp(powf(px[2], gamma) * 255.0f, 0.0f, 255.0f);
        float brightness = 0.2126f * px[0] + 0.7152f * px[1] + 0.0722f * px[2];
        *ci = rust_min(powf(brightness, gamma) * 67.0f, 67.0f);
    }
}

/* one pixel -> at most 26 bytes; returns the count */
static size_t put_pixel(uint8_t *dst, const float *px, int mode) {
    float fr, fg, fb, fci;
    pre_cast(px, mode, &fr, &fg, &fb, &fci);
    size_t n = 0;
    memcpy(dst, "\x1b[38;2;", 7); n = 7;
    n += put_u8(dst + n, rust_as_u8(fr)); dst[n++] = ';';
    n += put_u8(dst + n, rust_as_u8(fg)); dst[n++] = ';';
    n += put_u8(dst + n, rust_as_u8(fb)); dst[n++] = 'm';
    if (mode == 1) { dst[n++] = 0xE2; dst[n++] = 0x96; dst[n++] = 0x88; }     /* U+2588 */
    else dst[n++] = (uint8_t)RAMP[rust_as_usize(fci)];
    memcpy(dst + n, "\x1b[0m", 4); n += 4;
    return n;
}

size_t oracle_quantise(const float *pixels4, uint32_t w, uint32_t h, int mode, int threads,
                       uint8_t *out, size_t cap, uint32_t *row_lengths) {
    /* rows are built independently (lib.rs:503 into_par_iter) then concatenated (lib.rs:532) */
    size_t row_cap = (size_t)w * 26u + 2u;
    uint8_t *rows = (uint8_t *)malloc(row_cap * (h ? h : 1));
    uint32_t *lens = (uint32_t *)malloc(sizeof(uint32_t) * (h ? h : 1));
    if (!rows || !lens) { free(rows); free(lens); return 0; }
#ifdef _OPENMP
    int nt = threads > 0 ? threads : omp_get_max_threads();
#else
    (void)threads;
#endif
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt)
    for (int64_t j = 0; j < (int64_t)h; j++) {
        uint8_t *row = rows + row_cap * (size_t)j;
        size_t n = 0;
        for (uint32_t i = 0; i < w; i++) n += put_pixel(row + n, pixels4 + 4 * ((size_t)j * w + i), mode);
        row[n++] = '\r'; row[n++] = '\n';                                       /* lib.rs:527 */
        lens[j] = (uint32_t)n;
    }
    size_t total = 0;
    for (uint32_t j = 0; j < h; j++) {
        if (out && total + lens[j] <= cap) memcpy(out + total, rows + row_cap * (size_t)j, lens[j]);
        total += lens[j];
        if (row_lengths) row_lengths[j] = lens[j];
    }
    free(rows); free(lens);
    return total;
}

uint64_t oracle_quantise_boundary_count(const float *pixels4, uint32_t w, uint32_t h, int mode, float eps) {
    uint64_t n = 0;
    for (size_t k = 0; k < (size_t)w * h; k++) {
        float v[4];
        pre_cast(pixels4 + 4 * k, mode, &v[0], &v[1], &v[2], &v[3]);
        int near = 0;
        for (int c = 0; c < (mode == 1 ? 3 : 4); c++) {
            if (isnan(v[c]) || v[c] >= (c == 3 ? 67.0f : 255.0f)) continue;   /* saturated: no boundary */
            float f = v[c] - floorf(v[c]);
            if (v[c] > 0.0f && (f < eps || f > 1.0f - eps)) near = 1;
        }
        n += (uint64_t)near;
    }
    return n;
}
