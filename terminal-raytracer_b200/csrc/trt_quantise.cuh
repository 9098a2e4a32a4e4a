// trt_quantise.cuh -- on-device terminal quantiser: float4 frame -> ANSI byte stream.
//
// Replaces the reference's CPU row formatter /root/reference/src/lib.rs:503-532 (rayon, one String
// per row, concat).  Byte grammar per pixel (lib.rs:513,524):  ESC[38;2;R;G;Bm <glyph> ESC[0m,
// glyph = U+2588 (E2 96 88) in truecolor mode or one of the 68 ramp characters in ASCII mode;
// "\r\n" after each row (lib.rs:527).  Only the byte stream crosses PCIe.
//
// Three launches (HBM-bound; DESIGN.md "Quantiser"):
//   quant_pack   : 16 B/px in  -> 4 B/px packed (R,G,B,glyph) + per-block byte counts
//   quant_scan   : exclusive scan of the block counts (one CTA), total length
//   quant_emit   : 4 B/px in -> bytes staged in shared memory -> 16-byte aligned coalesced stores
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace trt {

#define TRT_Q_BLOCK 256            // pixels per CTA in pack/emit
#define TRT_Q_STAGE (TRT_Q_BLOCK * 28 + 32)

struct QuantParams {
    const float4 *src;             // frame (float4 per pixel)
    const uint64_t *row_src;       // per global row: element offset of its first pixel in src, or NULL (y*w)
    uint32_t width, height;
    int mode;                      // 0 ASCII, 1 truecolor
    uint32_t *packed;              // n_pixels
    uint32_t *block_len;           // n_blocks
    uint64_t *block_off;           // n_blocks + 1 (last = total)
    uint32_t n_blocks;
    uint8_t *out;
    uint64_t out_cap;
    // one-pass form (quant_fused_kernel)
    unsigned int *ticket;          // next tile to hand out (zeroed before the launch)
    unsigned long long *tile_state;// per tile: flag << 62 | byte count (aggregate of the tile, or inclusive prefix); zeroed
    uint64_t *row_start;           // rows + 1: stream offset of every row's first byte, [rows] = stream length
    uint64_t *total;               // stream length (device word the host copies)
    uint32_t n_tiles;
};

// lib.rs:521 -- 68 glyphs
__constant__ char c_ramp[69] = " .`^\",:;Il!i><~+_-?][}{1)(|\\tfjrxnuvczXYUJCLQ0OZmwqpdbkhao*#MW&8%B@$";

// Rust: (v).clamp(0.0, 255.0) as u8 -- clamp keeps NaN, `as u8` truncates, saturates, NaN -> 0
__device__ __forceinline__ uint32_t to_u8(float v) {
    if (!(v > 0.0f)) return 0u;
    if (v >= 255.0f) return 255u;
    return (uint32_t)v;
}
// Rust f32::powf(x, 0.3) = glibc powf on Linux.  Evaluated in double and rounded once: equals the
// correctly rounded f32 power except on double-rounding ties (byte-level differences
// against glibc powf are confined to quantisation-boundary pixels, counted by the tests).
__device__ __forceinline__ float powf_03(float x) { return (float)pow((double)x, (double)0.3f); }

__device__ __forceinline__ uint32_t quantise_pixel(float4 p, int mode) {
    uint32_t r, g, b, ci = 0;
    if (mode == 1) {                                    // lib.rs:510-512
        r = to_u8(sqrtf(p.x) * 255.0f);
        g = to_u8(sqrtf(p.y) * 255.0f);
        b = to_u8(sqrtf(p.z) * 255.0f);
    } else {                                            // lib.rs:515-523
        r = to_u8(powf_03(p.x) * 255.0f);
        g = to_u8(powf_03(p.y) * 255.0f);
        b = to_u8(powf_03(p.z) * 255.0f);
        float brightness = 0.2126f * p.x + 0.7152f * p.y + 0.0722f * p.z;
        float c = powf_03(brightness) * 67.0f;
        // f32::min(c, 67.0): NaN -> 67; then `as usize` (negative/NaN -> 0)
        if (c != c || c > 67.0f) c = 67.0f;
        ci = (c > 0.0f) ? (uint32_t)c : 0u;
    }
    return r | (g << 8) | (b << 16) | (ci << 24);
}

__device__ __forceinline__ uint32_t ndigits(uint32_t v) { return 1u + (v >= 10u) + (v >= 100u); }

__device__ __forceinline__ uint32_t pixel_len(uint32_t pk, int mode, bool row_end) {
    uint32_t n = 14u + ndigits(pk & 255u) + ndigits((pk >> 8) & 255u) + ndigits((pk >> 16) & 255u) + (mode == 1 ? 3u : 1u);
    return n + (row_end ? 2u : 0u);
}

__device__ __forceinline__ uint32_t put_u8(uint8_t *dst, uint32_t v) {
    uint32_t n = 0;
    if (v >= 100u) { dst[n++] = (uint8_t)('0' + v / 100u); v %= 100u; dst[n++] = (uint8_t)('0' + v / 10u); v %= 10u; }
    else if (v >= 10u) { dst[n++] = (uint8_t)('0' + v / 10u); v %= 10u; }
    dst[n++] = (uint8_t)('0' + v);
    return n;
}

__device__ __forceinline__ void put_pixel(uint8_t *d, uint32_t pk, int mode, bool row_end) {
    d[0] = 0x1b; d[1] = '['; d[2] = '3'; d[3] = '8'; d[4] = ';'; d[5] = '2'; d[6] = ';';
    uint32_t n = 7;
    n += put_u8(d + n, pk & 255u); d[n++] = ';';
    n += put_u8(d + n, (pk >> 8) & 255u); d[n++] = ';';
    n += put_u8(d + n, (pk >> 16) & 255u); d[n++] = 'm';
    if (mode == 1) { d[n++] = 0xE2; d[n++] = 0x96; d[n++] = 0x88; }
    else d[n++] = (uint8_t)c_ramp[pk >> 24];
    d[n++] = 0x1b; d[n++] = '['; d[n++] = '0'; d[n++] = 'm';
    if (row_end) { d[n++] = '\r'; d[n++] = '\n'; }
}

// CTA-wide exclusive scan of one u32 per thread (TRT_Q_BLOCK threads); returns the CTA total
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t &total) {
    __shared__ uint32_t warp_sums[TRT_Q_BLOCK / 32];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    uint32_t inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (uint32_t)o) inc += t;
    }
    if (lane == 31u) warp_sums[wid] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
    for (uint32_t w = 0; w < TRT_Q_BLOCK / 32; w++) {
        uint32_t s = warp_sums[w];
        if (w < wid) base += s;
        tot += s;
    }
    total = tot;
    __syncthreads();
    return base + inc - v;
}

__global__ void __launch_bounds__(TRT_Q_BLOCK) quant_pack_kernel(QuantParams q) {
    const uint64_t n = (uint64_t)q.width * q.height;
    const uint64_t i = (uint64_t)blockIdx.x * TRT_Q_BLOCK + threadIdx.x;
    uint32_t len = 0;
    if (i < n) {
        const uint32_t y = (uint32_t)(i / q.width), x = (uint32_t)(i - (uint64_t)y * q.width);
        const uint64_t s = q.row_src ? q.row_src[y] + x : i;
        const uint32_t pk = quantise_pixel(__ldg(&q.src[s]), q.mode);
        q.packed[i] = pk;
        len = pixel_len(pk, q.mode, x == q.width - 1u);
    }
    uint32_t total;
    block_exclusive_scan(len, total);
    if (threadIdx.x == 0) q.block_len[blockIdx.x] = total;
}

// one CTA of 1024 threads: exclusive scan of n_blocks counts into 64-bit offsets
__global__ void __launch_bounds__(1024) quant_scan_kernel(QuantParams q) {
    __shared__ uint64_t warp_sums[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    for (uint32_t base = 0; base < q.n_blocks; base += 1024u) {
        const uint32_t i = base + threadIdx.x;
        const uint64_t v = (i < q.n_blocks) ? q.block_len[i] : 0ull;
        uint64_t inc = v;
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        if (lane == 31u) warp_sums[wid] = inc;
        __syncthreads();
        uint64_t wbase = 0, tot = 0;
        for (uint32_t w = 0; w < 32u; w++) {
            uint64_t s = warp_sums[w];
            if (w < wid) wbase += s;
            tot += s;
        }
        const uint64_t carry = carry_s;
        if (i < q.n_blocks) q.block_off[i] = carry + wbase + inc - v;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) q.block_off[q.n_blocks] = carry_s;
}

__global__ void __launch_bounds__(TRT_Q_BLOCK) quant_emit_kernel(QuantParams q) {
    __shared__ __align__(16) uint8_t stage[TRT_Q_STAGE];
    const uint64_t n = (uint64_t)q.width * q.height;
    const uint64_t i = (uint64_t)blockIdx.x * TRT_Q_BLOCK + threadIdx.x;
    const uint64_t off = q.block_off[blockIdx.x];
    const uint32_t pad = (uint32_t)((uintptr_t)(q.out + off) & 15u);
    uint32_t pk = 0, len = 0;
    bool row_end = false;
    if (i < n) {
        pk = q.packed[i];
        row_end = (uint32_t)(i % q.width) == q.width - 1u;
        len = pixel_len(pk, q.mode, row_end);
    }
    uint32_t total;
    const uint32_t local = block_exclusive_scan(len, total);
    if (off + total > q.out_cap) return;                 // caller checks the length first
    if (i < n) put_pixel(stage + pad + local, pk, q.mode, row_end);
    __syncthreads();
    uint8_t *dst = q.out + off;                          // dst + k  <->  stage[pad + k]
    const uint32_t head = min(total, (16u - pad) & 15u); // bytes before the first 16-byte boundary
    const uint32_t body = (total - head) & ~15u;
    for (uint32_t k = threadIdx.x; k < head; k += TRT_Q_BLOCK) dst[k] = stage[pad + k];
    const uint4 *s4 = reinterpret_cast<const uint4 *>(stage + pad + head);
    uint4 *d4 = reinterpret_cast<uint4 *>(dst + head);
    for (uint32_t k = threadIdx.x; k < body / 16u; k += TRT_Q_BLOCK) d4[k] = s4[k];
    for (uint32_t k = head + body + threadIdx.x; k < total; k += TRT_Q_BLOCK) dst[k] = stage[pad + k];
}

// ---- one pass: quantise + scan + emit fused (decoupled look-back) ------------------------------------------------
// The three-launch form reads the frame (16 B/px), writes and re-reads 4 B/px of packed colours twice and 12 B per 256
// pixels of scan state.  Here one CTA takes a tile of TRT_QF_TILE pixels (dynamic tile index, so every predecessor has
// started), quantises it in registers, publishes the tile's byte count, obtains its stream offset by looking back over
// its predecessors' published counts (a warp reads 32 of them at a time), stages the text in shared memory and writes
// it with 16-byte stores: HBM traffic = 16 B/px in + the stream out, the algorithmic minimum.  The thread that owns the
// first pixel of a row also records where that row starts (row-addressed streams need every row's length).
#ifndef TRT_QF_PER_THREAD
#define TRT_QF_PER_THREAD 4
#endif
// resident CTAs the register allocation must allow.  The kernel is bound by the latency of a tile's dependent phases (load,
// scan, format, look-back, copy-out), so tiles in flight count: 1 (62 registers, 4 CTAs) 0.234 ms on a 3840x2160 frame,
// 5 (48 registers) 0.203 ms, 6 (40 registers, spills) 0.200-0.207 ms, 7 (32 registers) 0.208-0.213 ms (profiles/r2_quant_fused_raw.txt)
#ifndef TRT_QF_MIN_CTAS
#define TRT_QF_MIN_CTAS 5
#endif
#define TRT_QF_TILE (TRT_Q_BLOCK * TRT_QF_PER_THREAD)
#define TRT_QF_STAGE (TRT_QF_TILE * 28 + 32)
#define TRT_QF_FLAG_AGG (1ull << 62)
#define TRT_QF_FLAG_PFX (2ull << 62)
#define TRT_QF_VALUE (~(3ull << 62))

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// decimal text of a byte, left-aligned: char0 | char1 << 8 | char2 << 16 | digit count << 24 (one table entry per thread)
__device__ __forceinline__ uint32_t digits_entry(uint32_t v) {
    const uint32_t h = v / 100u, t = (v / 10u) % 10u, o = v % 10u;
    if (v >= 100u) return ('0' + h) | (('0' + t) << 8) | (('0' + o) << 16) | (3u << 24);
    if (v >= 10u) return ('0' + t) | (('0' + o) << 8) | (2u << 24);
    return ('0' + o) | (1u << 24);
}
// branch-free form of put_pixel over the digit table: three characters are always stored, the cursor advances by the
// digit count, the next store overwrites what was one too many
__device__ __forceinline__ void put_pixel_lut(uint8_t *d, uint32_t pk, int mode, bool row_end, const uint32_t *dig) {
    d[0] = 0x1b; d[1] = '['; d[2] = '3'; d[3] = '8'; d[4] = ';'; d[5] = '2'; d[6] = ';';
    uint32_t n = 7;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const uint32_t e = dig[(pk >> (8 * c)) & 255u];
        d[n] = (uint8_t)e; d[n + 1] = (uint8_t)(e >> 8); d[n + 2] = (uint8_t)(e >> 16);
        n += e >> 24;
        d[n++] = c == 2 ? 'm' : ';';
    }
    if (mode == 1) { d[n++] = 0xE2; d[n++] = 0x96; d[n++] = 0x88; }
    else d[n++] = (uint8_t)c_ramp[pk >> 24];
    d[n++] = 0x1b; d[n++] = '['; d[n++] = '0'; d[n++] = 'm';
    if (row_end) { d[n++] = '\r'; d[n++] = '\n'; }
}

__global__ void __launch_bounds__(TRT_Q_BLOCK, TRT_QF_MIN_CTAS) quant_fused_kernel(QuantParams q) {
    __shared__ __align__(16) uint8_t stage[TRT_QF_STAGE + 32];
    __shared__ uint32_t s_dig[256];
    __shared__ unsigned int s_tile;
    __shared__ unsigned long long s_base;
    if (threadIdx.x == 0) s_tile = atomicAdd(q.ticket, 1u);
    s_dig[threadIdx.x & 255u] = digits_entry(threadIdx.x & 255u);
    __syncthreads();
    const unsigned int tile = s_tile;
    const uint64_t n = (uint64_t)q.width * q.height;
    const uint64_t i0 = (uint64_t)tile * TRT_QF_TILE + (uint64_t)threadIdx.x * TRT_QF_PER_THREAD;
    uint32_t pk[TRT_QF_PER_THREAD], len[TRT_QF_PER_THREAD], mine = 0;
    uint32_t y = (uint32_t)(i0 / q.width), x = (uint32_t)(i0 - (uint64_t)y * q.width);
#pragma unroll
    for (int j = 0; j < TRT_QF_PER_THREAD; j++) {
        pk[j] = 0; len[j] = 0;
        if (i0 + j < n) {
            const uint64_t src = q.row_src ? q.row_src[y] + x : i0 + j;
            pk[j] = quantise_pixel(__ldg(&q.src[src]), q.mode);
            const uint32_t nd = (s_dig[pk[j] & 255u] >> 24) + (s_dig[(pk[j] >> 8) & 255u] >> 24) + (s_dig[(pk[j] >> 16) & 255u] >> 24);
            len[j] = (14u + nd + (q.mode == 1 ? 3u : 1u) + (x == q.width - 1u ? 2u : 0u)) | (x == 0u ? 0x80000000u : 0u);   // bit 31: first pixel of row y_j
            mine += len[j] & 0xffu;
            if (++x == q.width) { x = 0; y++; }
        }
    }
    uint32_t total;
    const uint32_t local = block_exclusive_scan(mine, total);
    // publish the tile's byte count at once: successors can look back over it while this tile is still formatting
    if (threadIdx.x == 0) st_release_u64(&q.tile_state[tile], (tile == 0 ? TRT_QF_FLAG_PFX : TRT_QF_FLAG_AGG) | total);
    // ---- format the tile's text at its LOCAL offsets (nothing here depends on where the tile lands in the stream) ----
    {
        uint32_t pos = local;
        uint32_t xx = (uint32_t)(i0 - (uint64_t)(i0 / q.width) * q.width);
#pragma unroll
        for (int j = 0; j < TRT_QF_PER_THREAD; j++) {
            if (i0 + j < n) {
                put_pixel_lut(stage + pos, pk[j], q.mode, xx == q.width - 1u, s_dig);
                pos += len[j] & 0xffu;
                if (++xx == q.width) xx = 0;
            }
        }
    }
    // ---- the tile's offset in the stream: decoupled look-back (warp 0, after its own share of the formatting) --------
    if (threadIdx.x < 32u) {
        const unsigned lane = threadIdx.x;
        unsigned long long excl = 0;
        int pred = (int)tile - 1;
        while (pred >= 0) {
            const int t = pred - (int)lane;
            unsigned long long st = TRT_QF_FLAG_PFX;                       // before the first tile: prefix 0
            if (t >= 0) {
                do { st = ld_acquire_u64(&q.tile_state[t]); } while ((st >> 62) == 0ull);
            }
            const unsigned has_pfx = __ballot_sync(0xffffffffu, (st >> 62) == 2ull);
            const int first = has_pfx ? __ffs(has_pfx) - 1 : 31;           // nearest predecessor that knows its prefix
            unsigned long long v = (int)lane <= first ? (st & TRT_QF_VALUE) : 0ull;
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            excl += v;
            if (has_pfx) break;
            pred -= 32;
        }
        if (lane == 0) {
            if (tile != 0) st_release_u64(&q.tile_state[tile], TRT_QF_FLAG_PFX | (excl + total));
            s_base = excl;
            if (tile == q.n_tiles - 1u) { *q.total = excl + total; if (q.row_start) q.row_start[q.height] = excl + total; }
        }
    }
    __syncthreads();                                     // the text is staged and the offset is known
    const uint64_t off = s_base;
    if (off + total > q.out_cap) return;                 // caller checks the length first
    if (q.row_start) {
        uint32_t pos = local;
        uint32_t yy = (uint32_t)(i0 / q.width);
#pragma unroll
        for (int j = 0; j < TRT_QF_PER_THREAD; j++) {
            if (i0 + j < n) {
                if (len[j] >> 31) q.row_start[yy] = off + pos;
                pos += len[j] & 0xffu;
            }
            // (row index of pixel j + 1: the row ends where a row-end pixel was just passed)
            if (i0 + j < n && (uint32_t)((i0 + j) - (uint64_t)yy * q.width) == q.width - 1u) yy++;
        }
    }
    // ---- stage[k] -> dst[k]: 16-byte stores to the 16-byte aligned part of the destination; the staged text sits at an
    // arbitrary byte phase against it, so every 16 bytes are assembled from five words with funnel shifts ---------------
    uint8_t *dst = q.out + off;
    const uint32_t head = min(total, (16u - (uint32_t)((uintptr_t)dst & 15u)) & 15u);     // bytes before the first 16-byte boundary
    const uint32_t body = (total - head) & ~15u;
    for (uint32_t k = threadIdx.x; k < head; k += TRT_Q_BLOCK) dst[k] = stage[k];
    const uint32_t sh = (head & 3u) * 8u;
    const uint32_t *sw = reinterpret_cast<const uint32_t *>(stage + (head & ~3u));
    uint4 *d4 = reinterpret_cast<uint4 *>(dst + head);
    for (uint32_t k = threadIdx.x; k < body / 16u; k += TRT_Q_BLOCK) {
        const uint32_t w0 = sw[4u * k], w1 = sw[4u * k + 1u], w2 = sw[4u * k + 2u], w3 = sw[4u * k + 3u], w4 = sw[4u * k + 4u];
        d4[k] = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
    }
    for (uint32_t k = head + body + threadIdx.x; k < total; k += TRT_Q_BLOCK) dst[k] = stage[k];
}

// row lengths from the row starts the fused kernel recorded
__global__ void quant_row_len_from_starts_kernel(const uint64_t *row_start, uint32_t rows, uint32_t *row_len) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) row_len[r] = (uint32_t)(row_start[r + 1] - row_start[r]);
}

// Byte length of every row of the last quantised frame (one warp per row): what a shard publishes so that the rows of
// all shards can be placed in one stream without moving a pixel (row-addressed streams, trt_api.cu).
__global__ void __launch_bounds__(256) quant_row_len_kernel(const uint32_t *packed, uint32_t width, uint32_t rows, int mode,
                                                            uint32_t *row_len) {
    const uint32_t r = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    if (r >= rows) return;
    const uint32_t *p = packed + (size_t)r * width;
    uint32_t n = 0;
    for (uint32_t x = lane; x < width; x += 32u) n += pixel_len(__ldg(&p[x]), mode, false);
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (lane == 0) row_len[r] = n + 2u;                  // "\r\n" (lib.rs:527)
}

}  // namespace trt
