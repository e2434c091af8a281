// quade_compact.cuh -- K5 compact (included by quade_kernels.cu)
//
// One warp per 32 consecutive pairs.  Per pair: the raw spans [name_off, qual_off+len) of its R1 and R2 records
// are prefetched one pair ahead with cp.async into a shared-memory ring; each record is then assembled as
//     '@' name ':' index [':' molecular] '\n' seq '\n' '+' '\n' qual '\n'      (FastqWriter.py:61-69,
//                                                                               FastqSeq.py:76-82)
// in a staging buffer that has the 16-byte alignment of its destination, and stored with aligned 128-bit
// vectors (first / last partial vector by bytes).  Qualities round-trip unchanged (chr(ord(x)-33+33)).
//
// Fast path (the record has no trailing blanks and a one-byte separator line, i.e. qual_off-seq_off == len+3):
// two funnel-shift passes over whole 16-byte vectors -- name, then seq"\n+\n"qual in one go -- that may spill
// garbage over the neighbouring bytes, then a byte phase that lays the suffix, '@', '+', the final '\n' and
// (when the name ends in the vector where the sequence starts) the clobbered name tail on top.
// General path (CRLF, decorated '+' line, trailing blanks): three exact segment copies with byte edges.
// Oversized records (span > 1 KiB): bytes straight from global memory.
#pragma once

namespace quade {
namespace {

constexpr int KT = 256;
constexpr int KCAPV = 64;                          // 16-byte vectors per staged span (1 KiB)
constexpr int KS_STRIDE = 16 + KCAPV * 16 + 32;    // slack vector before, data, two slack vectors after
constexpr int KO_VEC = KCAPV + 18;                 // the suffix can make the output longer than the span
constexpr int KO_BYTES = KO_VEC * 16;
constexpr int KBUF = 4;                            // 2 pairs x 2 reads

// 16 bytes starting at byte r (0..15) of the 32 bytes {A,B}; r is warp-uniform
__device__ __forceinline__ uint4 shift_bytes(const uint4 A, const uint4 B, uint32_t r)
{
    const uint32_t sh = (r & 3u) * 8u;
    uint32_t w0, w1, w2, w3, w4;
    switch (r >> 2) {
    case 0: w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; break;
    case 1: w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; break;
    case 2: w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; break;
    default: w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; break;
    }
    return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                      __funnelshift_r(w3, w4, sh));
}

// whole vectors [v0, v1) of O <- S shifted so that O byte d0 is S byte so (bytes outside the segment spill over)
__device__ __forceinline__ void copy_vectors(uint32_t S, uint32_t O, uint32_t v0, uint32_t v1, int delta, int lane)
{
    const uint32_t r = (uint32_t)delta & 15u;
#pragma unroll 1
    for (uint32_t v = v0 + lane; v < v1; v += 32) {
        const uint32_t sa = S + (uint32_t)((delta + (int)(16u * v)) & ~15);   // may reach into the slack vector before S
        sts128(O + 16u * v, shift_bytes(lds128(sa), lds128(sa + 16u), r));
    }
}

// exact copy of `len` bytes S[so ..) -> O[d0 ..): full vectors by funnel shift, the edges by bytes
__device__ __forceinline__ void copy_segment(uint32_t S, uint32_t O, uint32_t d0, uint32_t so, uint32_t len, int lane)
{
    const uint32_t d1 = d0 + len, fv = (d0 + 15u) >> 4, lv = d1 >> 4;
    if (lv > fv) copy_vectors(S, O, fv, lv, (int)so - (int)d0, lane);
    const uint32_t head_end = min(16u * fv, d1), t = lane & 15;
    if (lane < 16) {
        if (d0 + t < head_end) sts8(O + d0 + t, lds8(S + so + t));
    } else if (lv >= fv) {
        const uint32_t x = 16u * lv + t;
        if (x < d1) sts8(O + x, lds8(S + so + (x - d0)));
    }
}

// per-record metadata as it travels through the shuffles: name_off, out offset, name_len | (seq_off-name_off) << 16,
// seq_len | (qual_off-seq_off-seq_len) << 24.  A record that does not fit the packing is marked wide (0xff).
struct RecMeta {
    uint32_t name_off, ooff, m2, m3;
};
__device__ __forceinline__ RecMeta pack_meta(uint4 rc, uint32_t nlen, uint32_t ooff)
{
    const uint32_t gap1 = rc.y - rc.x, gap2 = rc.z - rc.y - rc.w;
    RecMeta m;
    m.name_off = rc.x;
    m.ooff = ooff;
    const bool wide = nlen > 0xffffu || gap1 > 0xffffu || rc.w > 0xffffffu || gap2 >= 0xffu;
    m.m2 = wide ? 0u : (nlen | gap1 << 16);
    m.m3 = wide ? 0xff000000u : (rc.w | gap2 << 24);
    return m;
}
__device__ __forceinline__ RecMeta shfl_meta(const RecMeta &m, int src)
{
    RecMeta r;
    r.name_off = __shfl_sync(0xffffffffu, m.name_off, src);
    r.ooff = __shfl_sync(0xffffffffu, m.ooff, src);
    r.m2 = __shfl_sync(0xffffffffu, m.m2, src);
    r.m3 = __shfl_sync(0xffffffffu, m.m3, src);
    return r;
}

} // namespace
} // namespace quade

#include "quade_compact2.cuh"
