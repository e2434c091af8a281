// quade_synth.cpp -- counter-based synthetic paired-end FASTQ generator.
//
// Test and benchmark utility (include/quade_synth.h), built into its own libquade_synth.so: it is NOT part of the
// demultiplexing path and libquade_b200.so does not contain it.  Every field of pair i is a pure function of
// (seed, i), so any sub-range of a data set can be produced independently (per chunk, per rank, per thread, per
// GPU thread) and is reproducible byte for byte: the same inline code fills host buffers (quade_synth_fill) and,
// compiled for the device, HBM-resident buffers (quade_synth_fill_device).
//
//   name   NB200:1:HQUADEB200:{lane 1-8}:{tile 1101-2678}:{x 1000-31999}:{y 1000-99999}   (shared by the mates)
//   descr  " {r}:N:0:0", r = 1 (R1), 4 (R2), 2 (I1), 3 (I2)
//   insert reads: read_len bases, uniform ACGT, N with p = 1/1024; phred 2/12/23/37 with p = 1/128, 5/128,
//          13/128, 109/128; bare '+' line
//   index reads: barcode (index_len) + molecular (mol_len) bases.  With probability `undetermined` the
//          barcodes are uniform random, else those of one uniformly chosen sample; molecular bases uniform;
//          phred per base uniform in [floor, 40], per-pair floor = 30 w.p. 0.9 else 2 (floor_mode 0) or
//          uniform in 2..40 (floor_mode 1)
#include "../../include/quade_synth.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#define HD __host__ __device__

namespace {

HD inline uint64_t mix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct Rng {
    uint64_t s;
    HD explicit Rng(uint64_t seed) : s(seed) {}
    HD uint64_t next() { return mix64(s++ * 0xD1342543DE82EF95ull + 0x632BE59BD9B4E019ull); }
    HD uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
};

struct PairFields {
    uint32_t lane, tile, x, y;
};

HD inline PairFields pair_fields(uint64_t seed, uint64_t i)
{
    Rng r(mix64(seed ^ mix64(i * 8 + 7)));
    PairFields f;
    f.lane = 1 + r.below(8);
    f.tile = 1101 + r.below(2678 - 1101 + 1);
    f.x = 1000 + r.below(31000);
    f.y = 1000 + r.below(99000);
    return f;
}

HD inline int ndigits(uint32_t v) { return v >= 10000 ? 5 : v >= 1000 ? 4 : v >= 100 ? 3 : v >= 10 ? 2 : 1; }

HD inline uint8_t *put_uint(uint8_t *w, uint32_t v)
{
    char tmp[12];
    int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n) *w++ = (uint8_t)tmp[--n];
    return w;
}

constexpr int kPrefixLen = 20;                 // "@NB200:1:HQUADEB200:"

HD inline uint32_t header_len(const PairFields &f) // '@' name ' ' r ":N:0:0" '\n'
{
    return kPrefixLen + 1 + 1 + ndigits(f.tile) + 1 + ndigits(f.x) + 1 + ndigits(f.y) + 8 + 1;
}

HD inline uint8_t *put_header(uint8_t *w, const PairFields &f, char r)
{
    const char prefix[kPrefixLen + 1] = "@NB200:1:HQUADEB200:";
    for (int k = 0; k < kPrefixLen; ++k) *w++ = (uint8_t)prefix[k];
    w = put_uint(w, f.lane); *w++ = ':';
    w = put_uint(w, f.tile); *w++ = ':';
    w = put_uint(w, f.x); *w++ = ':';
    w = put_uint(w, f.y);
    *w++ = ' '; *w++ = (uint8_t)r;
    const char tail[8] = ":N:0:0\n";
    for (int k = 0; k < 7; ++k) *w++ = (uint8_t)tail[k];
    return w;
}

HD inline uint8_t *put_insert(uint8_t *w, uint64_t seed, uint64_t i, int stream, int len)
{
    const char B[4] = {'A', 'C', 'G', 'T'};
    Rng r(mix64(seed ^ mix64(i * 8 + (uint64_t)stream)));
    for (int k = 0; k < len;) {
        uint64_t v = r.next();
        for (int j = 0; j < 5 && k < len; ++j, ++k, v >>= 12) *w++ = (v & 0xffc) == 0 ? 'N' : (uint8_t)B[v & 3];
    }
    *w++ = '\n'; *w++ = '+'; *w++ = '\n';
    for (int k = 0; k < len;) {
        uint64_t v = r.next();
        for (int j = 0; j < 9 && k < len; ++j, ++k, v >>= 7) {
            const uint32_t t = (uint32_t)(v & 127);
            *w++ = (uint8_t)(33 + (t < 1 ? 2 : t < 6 ? 12 : t < 19 ? 23 : 37));
        }
    }
    *w++ = '\n';
    return w;
}

// index reads of pair i: both are decided by one generator so that a sample's two barcodes go together
HD inline void put_index(uint8_t *w1, uint8_t *w2, const quade_synth_params &p, const uint8_t *barcodes, uint64_t i)
{
    const char B[4] = {'A', 'C', 'G', 'T'};
    Rng r(mix64(p.seed ^ mix64(i * 8 + 2)));
    const int il = p.index_len, ml = p.mol_len, ni = p.dual ? 2 : 1;
    const bool random_bc = p.n_samples == 0 || (double)(r.next() >> 11) * (1.0 / 9007199254740992.0) < p.undetermined;
    const uint32_t smp = p.n_samples ? r.below((uint32_t)p.n_samples) : 0;
    uint32_t floor_q;
    if (p.floor_mode == 0) floor_q = r.below(10) ? 30 : 2;
    else floor_q = 2 + r.below(39);
    uint8_t *w[2] = {w1, w2};
    for (int k = 0; k < ni; ++k) {
        uint8_t *o = w[k];
        if (random_bc) for (int j = 0; j < il; ++j) *o++ = (uint8_t)B[r.below(4)];
        else { const uint8_t *bc = barcodes + ((size_t)smp * ni + k) * il; for (int j = 0; j < il; ++j) *o++ = bc[j]; }
        for (int j = 0; j < ml; ++j) *o++ = (uint8_t)B[r.below(4)];
        *o++ = '\n'; *o++ = '+'; *o++ = '\n';
        for (int j = 0; j < il + ml; ++j) *o++ = (uint8_t)(33 + floor_q + r.below(41 - floor_q));
        *o++ = '\n';
    }
}

constexpr uint64_t kBlock = 4096; // pairs per work item

struct Sizes { uint64_t b[4]; };

Sizes block_sizes(const quade_synth_params &p, uint64_t first, uint64_t count)
{
    Sizes s{{0, 0, 0, 0}};
    const uint64_t ins = 2ull * p.read_len + 4, idx = 2ull * (p.index_len + p.mol_len) + 4;
    for (uint64_t i = first; i < first + count; ++i) {
        const uint32_t h = header_len(pair_fields(p.seed, i));
        s.b[0] += h + ins; s.b[1] += h + ins; s.b[2] += h + idx;
        if (p.dual) s.b[3] += h + idx;
    }
    return s;
}

int check(const quade_synth_params *p)
{
    if (!p || p->read_len < 0 || p->index_len < 1 || p->index_len > 32 || p->mol_len < 0 || p->n_samples < 0) return QUADE_SYNTH_ERR_ARG;
    return QUADE_SYNTH_OK;
}

} // namespace

extern "C" int quade_synth_barcodes(const quade_synth_params *p, uint8_t *barcodes)
{
    if (check(p) != QUADE_SYNTH_OK || !barcodes) return QUADE_SYNTH_ERR_ARG;
    static const char B[4] = {'A', 'C', 'G', 'T'};
    const int il = p->index_len, ni = p->dual ? 2 : 1, S = p->n_samples;
    for (int k = 0; k < ni; ++k) {
        Rng r(mix64(p->seed ^ mix64(0xBA5C0DE5ull + k)));
        int min_dist = 3, rejected = 0;
        for (int s = 0; s < S;) {
            uint8_t cand[32];
            for (int j = 0; j < il; ++j) cand[j] = (uint8_t)B[r.below(4)];
            bool ok = true;
            for (int t = 0; t < s && ok; ++t) {
                const uint8_t *o = barcodes + ((size_t)t * ni + k) * il;
                int d = 0;
                for (int j = 0; j < il; ++j) d += o[j] != cand[j];
                if (d < min_dist) ok = false;
            }
            if (!ok) {
                if (++rejected > 20000 && min_dist > 1) { --min_dist; rejected = 0; } // short barcodes: relax
                continue;
            }
            memcpy(barcodes + ((size_t)s * ni + k) * il, cand, il);
            ++s;
        }
    }
    return QUADE_SYNTH_OK;
}

extern "C" int quade_synth_sizes(const quade_synth_params *p, const uint8_t *, uint64_t first, uint64_t count, uint64_t nbytes[4])
{
    if (check(p) != QUADE_SYNTH_OK || !nbytes) return QUADE_SYNTH_ERR_ARG;
    const Sizes s = block_sizes(*p, first, count);
    for (int k = 0; k < 4; ++k) nbytes[k] = s.b[k];
    return QUADE_SYNTH_OK;
}

extern "C" int quade_synth_fill(const quade_synth_params *p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                                uint8_t *const stream[4], int n_threads)
{
    if (check(p) != QUADE_SYNTH_OK || !stream || (p->n_samples && !barcodes)) return QUADE_SYNTH_ERR_ARG;
    const uint64_t nblk = (count + kBlock - 1) / kBlock;
    if (!nblk) return QUADE_SYNTH_OK;
    if (n_threads < 1) n_threads = 1;
    if ((uint64_t)n_threads > nblk) n_threads = (int)nblk;
    std::vector<Sizes> base(nblk + 1);
    auto run = [&](auto fn) {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t) th.emplace_back([&, t] { for (uint64_t b = t; b < nblk; b += n_threads) fn(b); });
        for (auto &x : th) x.join();
    };
    run([&](uint64_t b) {
        const uint64_t lo = b * kBlock, n = std::min(kBlock, count - lo);
        base[b + 1] = block_sizes(*p, first + lo, n);
    });
    base[0] = Sizes{{0, 0, 0, 0}};
    for (uint64_t b = 0; b < nblk; ++b)
        for (int k = 0; k < 4; ++k) base[b + 1].b[k] += base[b].b[k];
    run([&](uint64_t b) {
        const uint64_t lo = b * kBlock, n = std::min(kBlock, count - lo);
        uint8_t *w[4];
        for (int k = 0; k < 4; ++k) w[k] = stream[k] ? stream[k] + base[b].b[k] : nullptr;
        const uint64_t idx_rec = 2ull * (p->index_len + p->mol_len) + 4;
        for (uint64_t i = first + lo; i < first + lo + n; ++i) {
            const PairFields f = pair_fields(p->seed, i);
            w[0] = put_insert(put_header(w[0], f, '1'), p->seed, i, 0, p->read_len);
            w[1] = put_insert(put_header(w[1], f, '4'), p->seed, i, 1, p->read_len);
            w[2] = put_header(w[2], f, '2');
            uint8_t *i2 = nullptr;
            if (p->dual) { w[3] = put_header(w[3], f, '3'); i2 = w[3]; }
            put_index(w[2], i2, *p, barcodes, i);
            w[2] += idx_rec;
            if (p->dual) w[3] += idx_rec;
        }
    });
    return QUADE_SYNTH_OK;
}

// ---- device side: the same bytes, generated in HBM -----------------------------------------------------------
namespace {

constexpr int kDevBlock = 256;   // pairs per CTA

// header bytes of the pairs of every CTA
__global__ void __launch_bounds__(kDevBlock) k_synth_sizes(quade_synth_params p, uint64_t first, uint64_t count, uint32_t *block_sum)
{
    __shared__ uint32_t s_sum;
    if (threadIdx.x == 0) s_sum = 0u;
    __syncthreads();
    const uint64_t i = (uint64_t)blockIdx.x * kDevBlock + threadIdx.x;
    uint32_t h = i < count ? header_len(pair_fields(p.seed, first + i)) : 0u;
    for (int o = 16; o; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_sum, h);
    __syncthreads();
    if (threadIdx.x == 0) block_sum[blockIdx.x] = s_sum;
}

// exclusive scan of the CTA sums (one CTA; 64-bit running total)
__global__ void __launch_bounds__(1024) k_synth_scan(const uint32_t *block_sum, unsigned long long *block_base, uint32_t nblk)
{
    __shared__ unsigned long long s_warp[33];
    __shared__ unsigned long long s_run;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_run = 0ull;
    __syncthreads();
    for (uint32_t b0 = 0; b0 < nblk; b0 += 1024u) {
        const unsigned long long v = b0 + tid < nblk ? block_sum[b0 + tid] : 0ull;
        unsigned long long inc = v;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long x = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += x;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const unsigned long long w = s_warp[lane];
            unsigned long long wi = w;
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long x = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += x;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_warp[32] = wi;
        }
        __syncthreads();
        if (b0 + tid < nblk) block_base[b0 + tid] = s_run + s_warp[warp] + inc - v;
        __syncthreads();
        if (tid == 0) s_run += s_warp[32];
        __syncthreads();
    }
    if (tid == 0) block_base[nblk] = s_run;
}

__global__ void __launch_bounds__(kDevBlock) k_synth_fill(quade_synth_params p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                                                          const unsigned long long *block_base, uint8_t *s0, uint8_t *s1, uint8_t *s2, uint8_t *s3)
{
    __shared__ uint32_t s_warp[kDevBlock / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t i = (uint64_t)blockIdx.x * kDevBlock + tid;
    PairFields f{0, 0, 0, 0};
    uint32_t h = 0u;
    if (i < count) {
        f = pair_fields(p.seed, first + i);
        h = header_len(f);
    }
    uint32_t inc = h;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += x;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0u;
    for (int w = 0; w < warp; ++w) wbase += s_warp[w];
    if (i >= count) return;
    const uint64_t hoff = block_base[blockIdx.x] + wbase + inc - h;           // header bytes of the pairs before this one
    const uint64_t ins = 2ull * p.read_len + 4, idx = 2ull * (p.index_len + p.mol_len) + 4;
    uint8_t *w0 = s0 + hoff + i * ins, *w1 = s1 + hoff + i * ins, *w2 = s2 + hoff + i * idx, *w3 = p.dual ? s3 + hoff + i * idx : nullptr;
    put_insert(put_header(w0, f, '1'), p.seed, first + i, 0, p.read_len);
    put_insert(put_header(w1, f, '4'), p.seed, first + i, 1, p.read_len);
    w2 = put_header(w2, f, '2');
    if (p.dual) w3 = put_header(w3, f, '3');
    put_index(w2, w3, p, barcodes, first + i);
}

} // namespace

extern "C" int quade_synth_fill_device(const quade_synth_params *p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                                       uint8_t *const d_stream[4], uint64_t nbytes[4], void *cuda_stream)
{
    if (check(p) != QUADE_SYNTH_OK || !d_stream || !nbytes || (p->n_samples && !barcodes)) return QUADE_SYNTH_ERR_ARG;
    for (int k = 0; k < 4; ++k) nbytes[k] = 0;
    if (!count) return QUADE_SYNTH_OK;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const uint32_t nblk = (uint32_t)((count + kDevBlock - 1) / kDevBlock);
    const size_t nbar = (size_t)std::max(1, p->n_samples) * (p->dual ? 2 : 1) * p->index_len;
    uint32_t *d_sum = nullptr;
    unsigned long long *d_base = nullptr;
    uint8_t *d_bar = nullptr;
    unsigned long long total = 0;
    cudaError_t e = cudaMalloc((void **)&d_sum, (size_t)nblk * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_base, ((size_t)nblk + 1) * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d_bar, nbar);
    if (e == cudaSuccess && p->n_samples) e = cudaMemcpyAsync(d_bar, barcodes, nbar, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        k_synth_sizes<<<nblk, kDevBlock, 0, st>>>(*p, first, count, d_sum);
        k_synth_scan<<<1, 1024, 0, st>>>(d_sum, d_base, nblk);
        k_synth_fill<<<nblk, kDevBlock, 0, st>>>(*p, d_bar, first, count, d_base, d_stream[0], d_stream[1], d_stream[2], d_stream[3]);
        e = cudaMemcpyAsync(&total, d_base + nblk, 8, cudaMemcpyDeviceToHost, st);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    cudaFree(d_sum);
    cudaFree(d_base);
    cudaFree(d_bar);
    if (e != cudaSuccess) return QUADE_SYNTH_ERR_CUDA;
    const uint64_t ins = 2ull * p->read_len + 4, idx = 2ull * (p->index_len + p->mol_len) + 4;
    nbytes[0] = nbytes[1] = total + count * ins;
    nbytes[2] = total + count * idx;
    nbytes[3] = p->dual ? nbytes[2] : 0;
    return QUADE_SYNTH_OK;
}
