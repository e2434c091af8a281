// quade_synth.cpp -- counter-based synthetic paired-end FASTQ generator.
//
// Host-side utility exported through the C-ABI for bench.py and the tests; it is NOT part of the
// demultiplexing path.  Every field of pair i is a pure function of (seed, i), so any sub-range of a data
// set can be produced independently (per chunk, per rank, per thread) and is reproducible byte for byte.
//
//   name   NB200:1:HQUADEB200:{lane 1-8}:{tile 1101-2678}:{x 1000-31999}:{y 1000-99999}   (shared by the mates)
//   descr  " {r}:N:0:0", r = 1 (R1), 4 (R2), 2 (I1), 3 (I2)
//   insert reads: read_len bases, uniform ACGT, N with p = 1/1024; phred 2/12/23/37 with p = 1/128, 5/128,
//          13/128, 109/128; bare '+' line
//   index reads: barcode (index_len) + molecular (mol_len) bases.  With probability `undetermined` the
//          barcodes are uniform random, else those of one uniformly chosen sample; molecular bases uniform;
//          phred per base uniform in [floor, 40], per-pair floor = 30 w.p. 0.9 else 2 (floor_mode 0) or
//          uniform in 2..40 (floor_mode 1)
#include "../../include/quade_b200.h"

#include <cstring>
#include <thread>
#include <vector>

namespace {

inline uint64_t mix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() { return mix64(s++ * 0xD1342543DE82EF95ull + 0x632BE59BD9B4E019ull); }
    uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
};

struct PairFields {
    uint32_t lane, tile, x, y;
};

inline PairFields pair_fields(uint64_t seed, uint64_t i)
{
    Rng r(mix64(seed ^ mix64(i * 8 + 7)));
    PairFields f;
    f.lane = 1 + r.below(8);
    f.tile = 1101 + r.below(2678 - 1101 + 1);
    f.x = 1000 + r.below(31000);
    f.y = 1000 + r.below(99000);
    return f;
}

inline int ndigits(uint32_t v) { return v >= 10000 ? 5 : v >= 1000 ? 4 : v >= 100 ? 3 : v >= 10 ? 2 : 1; }

inline uint8_t *put_uint(uint8_t *w, uint32_t v)
{
    char tmp[12];
    int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n) *w++ = (uint8_t)tmp[--n];
    return w;
}

const char kPrefix[] = "@NB200:1:HQUADEB200:";
constexpr int kPrefixLen = sizeof(kPrefix) - 1;

inline uint32_t header_len(const PairFields &f) // '@' name ' ' r ":N:0:0" '\n'
{
    return kPrefixLen + 1 + 1 + ndigits(f.tile) + 1 + ndigits(f.x) + 1 + ndigits(f.y) + 8 + 1;
}

inline uint8_t *put_header(uint8_t *w, const PairFields &f, char r)
{
    memcpy(w, kPrefix, kPrefixLen); w += kPrefixLen;
    w = put_uint(w, f.lane); *w++ = ':';
    w = put_uint(w, f.tile); *w++ = ':';
    w = put_uint(w, f.x); *w++ = ':';
    w = put_uint(w, f.y);
    *w++ = ' '; *w++ = (uint8_t)r;
    memcpy(w, ":N:0:0\n", 7); w += 7;
    return w;
}

inline uint8_t *put_insert(uint8_t *w, uint64_t seed, uint64_t i, int stream, int len)
{
    static const char B[4] = {'A', 'C', 'G', 'T'};
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
inline void put_index(uint8_t *w1, uint8_t *w2, const quade_synth_params &p, const uint8_t *barcodes, uint64_t i)
{
    static const char B[4] = {'A', 'C', 'G', 'T'};
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
        else { memcpy(o, barcodes + ((size_t)smp * ni + k) * il, il); o += il; }
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
    if (!p || p->read_len < 0 || p->index_len < 1 || p->index_len > 32 || p->mol_len < 0 || p->n_samples < 0) return QUADE_ERR_ARG;
    return QUADE_OK;
}

} // namespace

extern "C" int quade_synth_barcodes(const quade_synth_params *p, uint8_t *barcodes)
{
    if (check(p) != QUADE_OK || !barcodes) return QUADE_ERR_ARG;
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
    return QUADE_OK;
}

extern "C" int quade_synth_sizes(const quade_synth_params *p, const uint8_t *, uint64_t first, uint64_t count, uint64_t nbytes[4])
{
    if (check(p) != QUADE_OK || !nbytes) return QUADE_ERR_ARG;
    const Sizes s = block_sizes(*p, first, count);
    for (int k = 0; k < 4; ++k) nbytes[k] = s.b[k];
    return QUADE_OK;
}

extern "C" int quade_synth_fill(const quade_synth_params *p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                                uint8_t *const stream[4], int n_threads)
{
    if (check(p) != QUADE_OK || !stream || (p->n_samples && !barcodes)) return QUADE_ERR_ARG;
    const uint64_t nblk = (count + kBlock - 1) / kBlock;
    if (!nblk) return QUADE_OK;
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
    return QUADE_OK;
}
