// quade_compact2.cuh -- K5 compact: record assembly and the kernel (included by quade_compact.cuh)
#pragma once

namespace quade {
namespace {

// S: shared address of the staged span (byte 0 = global byte name_off & ~15), O: shared address of the staging
// buffer (byte 0 = arena byte ooff & ~15)
__device__ __forceinline__ void emit_record(const BatchDev &b, int k, uint32_t p, const RecMeta &m, uint32_t cls,
                                            uint32_t S, uint32_t O, int lane)
{
    const uint32_t suf_len = (cls >> 16) & 0xffu;
    const uint8_t *sufp = b.suffix + (size_t)p * b.suf_cap;
    const uint8_t *data = b.st[k].data;
    uint8_t *gout = b.out[k];
    uint32_t name_off = m.name_off, nlen, seq_off, qual_off, L;
    const bool wide = (m.m3 >> 24) == 0xffu;
    if (wide) {
        const uint4 rc = b.st[k].rec[p];
        nlen = b.st[k].vnlen[p];
        seq_off = rc.y; qual_off = rc.z; L = rc.w;
    } else {
        nlen = m.m2 & 0xffffu;
        seq_off = name_off + (m.m2 >> 16);
        L = m.m3 & 0xffffffu;
        qual_off = seq_off + L + (m.m3 >> 24);
    }
    const uint32_t total = nlen + suf_len + 2u * L + 5u;
    const uint32_t D1 = 1u + nlen + suf_len, D2 = D1 + L + 3u;
    const uint32_t al = name_off & ~15u, a = m.ooff & 15u;
    const uint32_t nv_src = (qual_off + L - al + 15u) >> 4;

    if (wide || nv_src > (uint32_t)KCAPV) {
        // oversized record: byte-wise straight from global memory
#pragma unroll 1
        for (uint32_t x = lane; x < total; x += 32) {
            uint32_t ch;
            if (x == 0u) ch = '@';
            else if (x < 1u + nlen) ch = data[name_off + x - 1u];
            else if (x < D1) ch = sufp[x - 1u - nlen];
            else if (x < D1 + L) ch = data[seq_off + x - D1];
            else if (x < D2) ch = (x - D1 - L == 1u) ? '+' : '\n';
            else if (x < D2 + L) ch = data[qual_off + x - D2];
            else ch = '\n';
            gout[(size_t)m.ooff + x] = (uint8_t)ch;
        }
        return;
    }
    const uint32_t dn0 = a + 1u, dn1 = dn0 + nlen, db0 = a + D1, endb = a + total;
    const uint32_t so_n = name_off - al, so_b = seq_off - al;
    if (qual_off - seq_off == L + 3u) {
        // fast path: name vectors, then seq "\n?\n" qual vectors, then the bytes in between
        copy_vectors(S, O, dn0 >> 4, (dn1 + 15u) >> 4, (int)so_n - (int)dn0, lane);
        __syncwarp();
        copy_vectors(S, O, db0 >> 4, (endb + 14u) >> 4, (int)so_b - (int)db0, lane);
        __syncwarp();
#pragma unroll 1
        for (uint32_t t = lane; t < suf_len; t += 32) sts8(O + dn1 + t, sufp[t]);
        const uint32_t vb = db0 & ~15u;               // the sequence pass clobbered the name bytes of this vector
        if (dn1 > vb && lane >= 16) {
            const uint32_t x = vb + (lane - 16);
            if (x >= dn0 && x < dn1) sts8(O + x, lds8(S + so_n + (x - dn0)));
        }
        if (lane < 3) sts8(O + (lane == 0 ? a : lane == 1 ? db0 + L + 1u : endb - 1u), lane == 0 ? '@' : lane == 1 ? '+' : '\n');
    } else {
        copy_segment(S, O, dn0, so_n, nlen, lane);
        copy_segment(S, O, db0, so_b, L, lane);
        copy_segment(S, O, a + D2, qual_off - al, L, lane);
#pragma unroll 1
        for (uint32_t t = lane; t < suf_len; t += 32) sts8(O + dn1 + t, sufp[t]);
        if (lane == 0) sts8(O + a, '@');
        else if (lane <= 3) sts8(O + db0 + L + (lane - 1), lane == 2 ? '+' : '\n');
        else if (lane == 4) sts8(O + endb - 1u, '\n');
    }
    __syncwarp();

    // aligned 128-bit stores; first / last partial vector by bytes
    uint8_t *gdst = gout + (size_t)(m.ooff - a);
    const uint32_t F0 = a ? 1u : 0u, F1 = endb >> 4;
    uint32_t head_end, tail_beg;
    if (F1 >= F0) {
        uint4 *G4 = reinterpret_cast<uint4 *>(gdst);
#pragma unroll 1
        for (uint32_t v = F0 + lane; v < F1; v += 32) G4[v] = lds128(O + 16u * v);
        head_end = 16u * F0;
        tail_beg = 16u * F1;
    } else {
        head_end = endb;
        tail_beg = endb;
    }
    const uint32_t t = lane & 15;
    if (lane < 16) {
        if (a + t < head_end) gdst[a + t] = (uint8_t)lds8(O + a + t);
    } else {
        if (tail_beg + t < endb) gdst[tail_beg + t] = (uint8_t)lds8(O + tail_beg + t);
    }
    __syncwarp();
}

__global__ void __launch_bounds__(KT) k_compact(const __grid_constant__ BatchDev b)
{
    extern __shared__ __align__(16) uint8_t k_smem[];
    const BatchState *St = b.state;
    if (St->err_capacity) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t n_pairs = St->n_pairs, n_tasks = (n_pairs + 31u) >> 5;
    const uint32_t sbase = smem_u32(k_smem) + (uint32_t)warp * (KBUF * KS_STRIDE + KO_BYTES);
    const uint32_t O = sbase + KBUF * KS_STRIDE;

    for (uint32_t task = blockIdx.x * (KT / 32) + warp; task < n_tasks; task += gridDim.x * (KT / 32)) {
        const uint32_t p = task * 32u + lane;
        uint32_t cls = 0;
        RecMeta me[2] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
        if (p < n_pairs) {
            cls = b.cls[p];
            me[0] = pack_meta(b.st[0].rec[p], b.st[0].vnlen[p], b.out1[p]);
            me[1] = pack_meta(b.st[1].rec[p], b.st[1].vnlen[p], b.out2[p]);
        }
        const int np = (int)min(32u, n_pairs - task * 32u);

        // prefetch of the two raw spans of one pair into ring buffers buf0, buf0+1
        auto issue = [&](const RecMeta (&m)[2], uint32_t c, int buf0) {
            if (c >> 31) {
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if ((m[k].m3 >> 24) == 0xffu) continue;
                    const uint32_t al = m[k].name_off & ~15u;
                    const uint32_t s1 = m[k].name_off + (m[k].m2 >> 16) + 2u * (m[k].m3 & 0xffffffu) + (m[k].m3 >> 24);
                    const uint32_t nv = (s1 - al + 15u) >> 4;
                    if (nv > (uint32_t)KCAPV) continue;
                    const uint8_t *src = b.st[k].data + al;
                    const uint32_t dst = sbase + (uint32_t)(buf0 + k) * KS_STRIDE + 16u;
#pragma unroll 1
                    for (uint32_t v = lane; v < nv; v += 32) cp_async16(dst + 16u * v, src + 16u * v);
                }
            }
            cp_async_commit();
        };

        RecMeta cur[2], nxt[2];
        uint32_t ccls = __shfl_sync(0xffffffffu, cls, 0), ncls = 0;
        cur[0] = shfl_meta(me[0], 0);
        cur[1] = shfl_meta(me[1], 0);
        nxt[0] = cur[0];
        nxt[1] = cur[1];
        issue(cur, ccls, 0);
#pragma unroll 1
        for (int i = 0; i < np; ++i) {
            if (i + 1 < np) {
                ncls = __shfl_sync(0xffffffffu, cls, i + 1);
                nxt[0] = shfl_meta(me[0], i + 1);
                nxt[1] = shfl_meta(me[1], i + 1);
                issue(nxt, ncls, ((i + 1) & 1) * 2);
            } else {
                cp_async_commit();
            }
            cp_async_wait<1>();
            __syncwarp();
            if (ccls >> 31) {                                // else the bin's writer is off (Sample.py:74,82,90)
                const uint32_t S0 = sbase + (uint32_t)((i & 1) * 2) * KS_STRIDE + 16u;
                emit_record(b, 0, task * 32u + i, cur[0], ccls, S0, O, lane);
                emit_record(b, 1, task * 32u + i, cur[1], ccls, S0 + KS_STRIDE, O, lane);
            }
            cur[0] = nxt[0];
            cur[1] = nxt[1];
            ccls = ncls;
        }
        cp_async_wait<0>();
        __syncwarp();
    }
}

constexpr size_t kCompactSmem = (size_t)(KT / 32) * (KBUF * KS_STRIDE + KO_BYTES);

} // namespace
} // namespace quade
