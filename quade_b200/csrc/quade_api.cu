// quade_api.cu -- the C-ABI of libquade_b200.so (include/quade_b200.h): contexts, batch slots, the four-stream
// pipeline (H2D copy | kernels | small per-batch results | D2H copy) and the host-side perfect-hash builder for the sample registry.
#include "quade_dev.h"
#include "../../include/quade_b200.h"

#include <algorithm>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace quade;

static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

namespace quade {
void set_last_error(const char *text) { snprintf(g_err, sizeof(g_err), "%s", text); }
} // namespace quade

#define CU(call)                                                                                                   \
    do {                                                                                                           \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess) return fail(QUADE_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_));              \
    } while (0)

namespace {

constexpr int kEvents = QUADE_K_COUNT + 1;

struct Slot {
    uint8_t *arena = nullptr;     // one device allocation, carved below
    size_t arena_bytes = 0;
    uint8_t *d_in[4] = {nullptr, nullptr, nullptr, nullptr};
    uint8_t *zero_base = nullptr; // region cleared before every batch
    size_t zero_bytes = 0;
    BatchDev dev;
    // pinned host mirrors
    uint8_t *h_out[2] = {nullptr, nullptr};
    bool d2h_issued = false;      // the output copy of the batch in this slot has been enqueued (quade_peek / quade_collect)
    BatchState *h_state = nullptr;
    unsigned long long *h_bin_off = nullptr;
    long long *h_counters = nullptr;
    cudaEvent_t ev_h2d0, ev_h2d1, ev_k[kEvents], ev_done, ev_d2h0, ev_d2h1;
    bool busy = false, device_mode = false, has_h2d = false;
    uint64_t ticket = 0;
    uint64_t launches = 0;
};

} // namespace

struct quade_ctx {
    int device = 0;
    quade_params prm;
    int n_streams = 3, n_bins = 1, n_slots = 2;
    uint64_t cap_bytes = 0, cap_pairs = 0, cap_rec = 0;
    uint32_t suf_cap = 4, pairs_per_block = 4096, cap_blocks = 1;
    uint64_t out_cap = 0;
    cudaStream_t s_copy = nullptr, s_compute = nullptr, s_d2h = nullptr, s_back = nullptr, s_meta = nullptr;
    bool overlap = false;                // k_compact3 of batch i on s_back beside the framing of batch i + 1
    std::vector<Slot> slots;
    int head = 0, tail = 0, in_flight = 0;
    bool samples_set = false;
    uint4 *d_entries = nullptr;
    uint16_t *d_disp = nullptr;
    HashTable table;
    std::vector<long long> cumulative;
};

static inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" int quade_version(void) { return QUADE_ABI_VERSION; }
extern "C" const char *quade_last_error(void) { return g_err; }

extern "C" int quade_device_count(int *count)
{
    if (!count) return fail(QUADE_ERR_ARG, "count is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        return fail(QUADE_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    }
    *count = n;
    return QUADE_OK;
}

extern "C" int quade_host_alloc(void **ptr, uint64_t nbytes)
{
    if (!ptr) return fail(QUADE_ERR_ARG, "ptr is NULL");
    CU(cudaHostAlloc(ptr, nbytes ? nbytes : 1, cudaHostAllocDefault));
    return QUADE_OK;
}

extern "C" int quade_host_free(void *ptr)
{
    if (ptr) CU(cudaFreeHost(ptr));
    return QUADE_OK;
}

static int slot_alloc(quade_ctx *c, Slot &sl)
{
    const int ns = c->n_streams, nb = c->n_bins;
    const size_t cap_rec = c->cap_rec, cap_pairs = c->cap_pairs;
    const size_t tiles = c->cap_bytes / kFrameTile + 2;
    // sizes
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += up256(bytes); return o; };
    size_t o_in[4], o_rec[4], o_nv[4], o_sfx[4], o_loc[4], o_vraw[4], o_tt[4];
    uint32_t sfx_slot[4] = {0, 0, 0, 0};
    const size_t o_zero = take(0);
    const size_t o_state = take(sizeof(BatchState));
    const size_t o_counters = take((size_t)nb * 8);
    const size_t zero_end = off;
    for (int s = 0; s < ns; ++s) {
        o_tt[s] = take(tiles * 32);
        o_in[s] = take(c->cap_bytes + 256);
        o_rec[s] = take(cap_rec * 16);
        o_nv[s] = take(cap_rec * 8);
        o_loc[s] = take(cap_rec * 4);
        o_vraw[s] = take(cap_rec * 4);
        o_sfx[s] = 0;
        if (s >= 2) {   // index reads: raw span bytes for the name suffix
            const uint32_t wi = (uint32_t)(c->prm.span[s - 2][1] - c->prm.span[s - 2][0]);
            const uint32_t wm = (uint32_t)(c->prm.span[s][1] - c->prm.span[s][0]);
            sfx_slot[s] = (((wi + 3u) & ~3u) + ((wm + 3u) & ~3u) + 15u) & ~15u;
            if (sfx_slot[s] < 16u) sfx_slot[s] = 16u;
            o_sfx[s] = take(cap_rec * sfx_slot[s] + 64);
        }
    }
    const size_t redo_cap = (size_t)ns * tiles;
    const size_t o_redo = take(redo_cap * 4);
    const size_t o_cls = take(cap_pairs * 4), o_len1 = take(cap_pairs * 4), o_len2 = take(cap_pairs * 4);
    const size_t o_out1 = take(cap_pairs * 4), o_out2 = take(cap_pairs * 4);
    const size_t o_suf = take(cap_pairs * c->suf_cap + 64);
    const size_t o_hist = take((size_t)c->cap_blocks * 2 * nb * 4);
    const size_t o_btot = take((size_t)2 * nb * 8), o_boff = take((size_t)2 * (nb + 1) * 8);
    const size_t o_o1 = take(c->out_cap + 256), o_o2 = take(c->out_cap + 256);

    cudaError_t e = cudaMalloc((void **)&sl.arena, off);
    if (e != cudaSuccess) return fail(QUADE_ERR_NOMEM, "cudaMalloc of %zu bytes for a batch slot: %s", off, cudaGetErrorString(e));
    sl.arena_bytes = off;
    uint8_t *A = sl.arena;
    sl.zero_base = A + o_zero;
    sl.zero_bytes = zero_end - o_zero;

    BatchDev &d = sl.dev;
    memset(&d, 0, sizeof(d));
    d.n_streams = ns;
    d.dual = c->prm.dual ? 1 : 0;
    for (int i = 0; i < 4; ++i) { d.span[i][0] = c->prm.span[i][0]; d.span[i][1] = c->prm.span[i][1]; }
    d.min_qual = c->prm.min_qual;
    d.wr_und = c->prm.write_undetermined ? 1 : 0;
    d.wr_pass = c->prm.write_pass ? 1 : 0;
    d.wr_fail = c->prm.write_fail ? 1 : 0;
    d.n_bins = nb;
    d.cap_pairs = (uint32_t)cap_pairs;
    d.pairs_per_block = c->pairs_per_block;
    d.cap_blocks = c->cap_blocks;
    d.suf_cap = c->suf_cap;
    d.out_cap[0] = d.out_cap[1] = c->out_cap;
    d.state = reinterpret_cast<BatchState *>(A + o_state);
    d.counters = reinterpret_cast<long long *>(A + o_counters);
    for (int s = 0; s < ns; ++s) {
        StreamDev &st = d.st[s];
        sl.d_in[s] = A + o_in[s];
        st.data = sl.d_in[s];
        st.cap_rec = (uint32_t)cap_rec;
        st.rec = reinterpret_cast<uint4 *>(A + o_rec[s]);
        st.nv = reinterpret_cast<uint2 *>(A + o_nv[s]);
        st.loc = reinterpret_cast<uint32_t *>(A + o_loc[s]);
        st.vraw = reinterpret_cast<uint32_t *>(A + o_vraw[s]);
        st.sfx = s >= 2 ? A + o_sfx[s] : nullptr;
        st.sfx_slot = sfx_slot[s];
        st.wi = s >= 2 ? (uint32_t)(c->prm.span[s - 2][1] - c->prm.span[s - 2][0]) : 0u;
        st.wm = s >= 2 ? (uint32_t)(c->prm.span[s][1] - c->prm.span[s][0]) : 0u;
        st.tile_tab = reinterpret_cast<uint4 *>(A + o_tt[s]);
    }
    d.redo_list = reinterpret_cast<uint32_t *>(A + o_redo);
    d.redo_cap = (uint32_t)redo_cap;
    d.cls = reinterpret_cast<uint32_t *>(A + o_cls);
    d.len1 = reinterpret_cast<uint32_t *>(A + o_len1);
    d.len2 = reinterpret_cast<uint32_t *>(A + o_len2);
    d.out1 = reinterpret_cast<uint32_t *>(A + o_out1);
    d.out2 = reinterpret_cast<uint32_t *>(A + o_out2);
    d.suffix = A + o_suf;
    d.hist = reinterpret_cast<uint32_t *>(A + o_hist);
    d.bin_total = reinterpret_cast<unsigned long long *>(A + o_btot);
    d.bin_off = reinterpret_cast<unsigned long long *>(A + o_boff);
    d.out[0] = A + o_o1;
    d.out[1] = A + o_o2;

    // pinned mirrors (the two output arenas only when asked for: device-mode callers never need them)
    CU(cudaHostAlloc((void **)&sl.h_state, sizeof(BatchState), cudaHostAllocDefault));
    CU(cudaHostAlloc((void **)&sl.h_bin_off, (size_t)2 * (nb + 1) * 8, cudaHostAllocDefault));
    CU(cudaHostAlloc((void **)&sl.h_counters, (size_t)nb * 8, cudaHostAllocDefault));
    if (!(c->prm.flags & QUADE_CTX_DEVICE_ONLY))
        for (int k = 0; k < 2; ++k) CU(cudaHostAlloc((void **)&sl.h_out[k], c->out_cap + 256, cudaHostAllocDefault));
    CU(cudaEventCreate(&sl.ev_h2d0));
    CU(cudaEventCreate(&sl.ev_h2d1));
    for (int i = 0; i < kEvents; ++i) CU(cudaEventCreate(&sl.ev_k[i]));
    CU(cudaEventCreate(&sl.ev_done));
    CU(cudaEventCreate(&sl.ev_d2h0));
    CU(cudaEventCreate(&sl.ev_d2h1));
    return QUADE_OK;
}

static void slot_free(Slot &sl)
{
    if (sl.arena) cudaFree(sl.arena);
    for (int k = 0; k < 2; ++k)
        if (sl.h_out[k]) cudaFreeHost(sl.h_out[k]);
    if (sl.h_state) {
        cudaFreeHost(sl.h_state);
        cudaFreeHost(sl.h_bin_off);
        cudaFreeHost(sl.h_counters);
        cudaEventDestroy(sl.ev_h2d0);
        cudaEventDestroy(sl.ev_h2d1);
        for (int i = 0; i < kEvents; ++i) cudaEventDestroy(sl.ev_k[i]);
        cudaEventDestroy(sl.ev_done);
        cudaEventDestroy(sl.ev_d2h0);
        cudaEventDestroy(sl.ev_d2h1);
    }
    sl = Slot();
}

extern "C" int quade_ctx_create(int device, const quade_params *p, quade_ctx **out)
{
    if (!p || !out) return fail(QUADE_ERR_ARG, "params/out is NULL");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(QUADE_ERR_CUDA, "no CUDA device: libquade_b200 has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(QUADE_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
    if (p->n_samples < 0 || p->n_samples > 8191) return fail(QUADE_ERR_UNSUPPORTED, "n_samples %d not in [0, 8191]", p->n_samples);
    uint32_t width = 0;
    for (int i = 0; i < 4; ++i) {
        if (p->span[i][0] < 0 || p->span[i][1] < p->span[i][0])
            return fail(QUADE_ERR_ARG, "span %d: need 0 <= start <= end (src/Quade.py:277-279)", i);
        width += (uint32_t)(p->span[i][1] - p->span[i][0]);
    }
    if (!p->dual && (p->span[1][1] || p->span[3][1])) return fail(QUADE_ERR_ARG, "index-2 spans given without dual index");
    if (width + 3 > 255) return fail(QUADE_ERR_UNSUPPORTED, "index + molecular spans wider than 252 bases");
    CU(cudaSetDevice(device));

    quade_ctx *c = new quade_ctx();
    c->device = device;
    c->prm = *p;
    c->n_streams = p->dual ? 4 : 3;
    c->n_bins = 1 + 2 * p->n_samples;
    c->n_slots = p->n_slots ? p->n_slots : 2;
    if (c->n_slots < 1 || c->n_slots > 8) { delete c; return fail(QUADE_ERR_ARG, "n_slots must be 1..8"); }
    c->cap_bytes = p->max_stream_bytes ? p->max_stream_bytes : (256ull << 20);
    c->cap_bytes = (c->cap_bytes + 255) & ~255ull;
    c->cap_pairs = p->max_pairs ? p->max_pairs : c->cap_bytes / 64;
    if (c->cap_pairs < 32) c->cap_pairs = 32;
    // records per file and batch the scratch arrays hold: by default enough for 24-byte records (an index file of
    // 4-base reads under short headers); a tile reserves one slot more than it may need
    c->cap_rec = (p->max_records ? p->max_records : c->cap_bytes / 24) + c->cap_bytes / kFrameTile + 1024;
    c->suf_cap = (width + 3 + 15) & ~15u;   // 16-byte slots: k_compact3 stages them with 128-bit loads
    c->out_cap = c->cap_bytes + c->cap_pairs * c->suf_cap;
    if (c->cap_bytes > kMaxStreamBytes || c->out_cap >= (1ull << 32) || c->cap_rec >= (1ull << 31)) {
        delete c;
        return fail(QUADE_ERR_ARG, "batch capacity too large: stream bytes and output arena must stay below 4 GiB");
    }
    uint32_t ppb = 4096;   // pairs per histogram block (1024 / 2048 / 4096: classify + bin scans + offsets 0.153 / 0.142 / 0.138 ms)
    if (const char *v = getenv("QUADE_PPB")) ppb = (uint32_t)std::max(256, atoi(v));
    while (ppb < (uint32_t)c->n_bins && ppb < 65536u) ppb <<= 1;
    c->pairs_per_block = ppb;
    c->cap_blocks = (uint32_t)((c->cap_pairs + ppb - 1) / ppb);
    c->cumulative.assign(c->n_bins, 0);
    if (const char *v = getenv("QUADE_OVERLAP")) c->overlap = atoi(v) != 0 && c->n_slots >= 2;
    int rc = QUADE_OK;
    if (cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->s_compute, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->s_back, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->s_meta, cudaStreamNonBlocking) != cudaSuccess)
        rc = fail(QUADE_ERR_CUDA, "cudaStreamCreate failed");
    c->slots.resize(c->n_slots);
    for (int i = 0; i < c->n_slots && rc == QUADE_OK; ++i) rc = slot_alloc(c, c->slots[i]);
    if (rc != QUADE_OK) {
        char keep[512];
        snprintf(keep, sizeof(keep), "%s", g_err);
        quade_ctx_destroy(c);
        snprintf(g_err, sizeof(g_err), "%s", keep);
        return rc;
    }
    *out = c;
    return QUADE_OK;
}

extern "C" int quade_ctx_destroy(quade_ctx *c)
{
    if (!c) return QUADE_OK;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
#if QUADE_CHECKS
    {
        unsigned int t[2];
        debug_check_tally(t);
        fprintf(stderr, "QUADE_CHECKS: %u violations (first code %u)\n", t[0], t[1]);
    }
#endif
    for (auto &sl : c->slots) slot_free(sl);
    if (c->d_entries) cudaFree(c->d_entries);
    if (c->d_disp) cudaFree(c->d_disp);
    if (c->s_copy) cudaStreamDestroy(c->s_copy);
    if (c->s_compute) cudaStreamDestroy(c->s_compute);
    if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
    if (c->s_back) cudaStreamDestroy(c->s_back);
    if (c->s_meta) cudaStreamDestroy(c->s_meta);
    delete c;
    return QUADE_OK;
}

// ---- sample registry: INDEX_TO_SAMPLE (src/Sample.py:37,151) as a hash-and-displace perfect hash -----------
namespace {
struct PackedKey {
    uint64_t key;
    uint32_t nmask, len;
};
} // namespace

extern "C" int quade_ctx_set_samples(quade_ctx *c, const uint8_t *keys, const uint32_t *key_off)
{
    if (!c) return fail(QUADE_ERR_ARG, "ctx is NULL");
    const int S = c->prm.n_samples;
    if (S > 0 && (!keys || !key_off)) return fail(QUADE_ERR_ARG, "keys/key_off is NULL");
    if (c->in_flight) return fail(QUADE_ERR_STATE, "batches in flight");
    CU(cudaSetDevice(c->device));
    std::vector<PackedKey> pk(S);
    for (int s = 0; s < S; ++s) {
        const uint8_t *k = keys + key_off[s];
        const uint32_t len = key_off[s + 1] - key_off[s];
        if (len == 0) return fail(QUADE_ERR_BAD_SAMPLE, "sample %d has an empty index", s);
        if (len > 32) return fail(QUADE_ERR_UNSUPPORTED, "sample %d: fused index of %u bases (max 32)", s, len);
        PackedKey q{0, 0, len};
        for (uint32_t i = 0; i < len; ++i) {
            const uint8_t ch = k[i];
            if (ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T') q.key |= (uint64_t)((ch >> 1) & 3u) << (2 * i);
            else if (ch == 'N') q.nmask |= 1u << i;
            else return fail(QUADE_ERR_BAD_SAMPLE, "sample %d: index is not a DNA sequence (ATCGN only, src/Sample.py:171-175)", s);
        }
        pk[s] = q;
    }
    for (int a = 0; a < S; ++a)  // Sample.py:140 uniqueness assert
        for (int b2 = a + 1; b2 < S; ++b2)
            if (pk[a].key == pk[b2].key && pk[a].nmask == pk[b2].nmask && pk[a].len == pk[b2].len)
                return fail(QUADE_ERR_BAD_SAMPLE, "samples %d and %d share the same index", a, b2);

    uint32_t n_slots = 16, n_buckets = 4;
    while (n_slots < 2u * (uint32_t)S) n_slots <<= 1;   // load <= 50 %: 384 samples -> 16 KiB of table per k_classify CTA, 8 CTAs per SM
    while (n_buckets < (uint32_t)(S + 1) / 2) n_buckets <<= 1;
    std::vector<uint4> entries;
    std::vector<uint16_t> disp;
    uint64_t seed = 0;
    bool ok = false;
    for (int attempt = 0; attempt < 256 && !ok; ++attempt) {
        seed = mix64(0x5EEDull + attempt);
        entries.assign(n_slots, make_uint4(0, 0, 0, 0));
        disp.assign(n_buckets, 0);
        std::vector<std::vector<int>> bucket(n_buckets);
        std::vector<uint64_t> h(S);
        for (int s = 0; s < S; ++s) {
            h[s] = key_hash(pk[s].key, pk[s].nmask, pk[s].len, seed);
            bucket[(uint32_t)h[s] & (n_buckets - 1)].push_back(s);
        }
        std::vector<int> order(n_buckets);
        for (uint32_t i = 0; i < n_buckets; ++i) order[i] = (int)i;
        std::sort(order.begin(), order.end(), [&](int x, int y) {
            return bucket[x].size() != bucket[y].size() ? bucket[x].size() > bucket[y].size() : x < y;
        });
        ok = true;
        for (int bi : order) {
            const auto &bk = bucket[bi];
            if (bk.empty()) continue;
            bool placed = false;
            for (uint32_t d = 0; d < 65536 && !placed; ++d) {
                std::vector<uint32_t> sl;
                bool clash = false;
                for (int s : bk) {
                    uint32_t x = key_slot(h[s], d, n_slots - 1);
                    if (entries[x].w || std::find(sl.begin(), sl.end(), x) != sl.end()) { clash = true; break; }
                    sl.push_back(x);
                }
                if (clash) continue;
                for (size_t i = 0; i < bk.size(); ++i) {
                    const PackedKey &q = pk[bk[i]];
                    entries[sl[i]] = make_uint4((uint32_t)q.key, (uint32_t)(q.key >> 32), q.nmask, q.len | (uint32_t)(bk[i] + 1) << 8);
                }
                disp[bi] = (uint16_t)d;
                placed = true;
            }
            if (!placed) { ok = false; break; }
        }
    }
    if (!ok) return fail(QUADE_ERR_BAD_SAMPLE, "could not build a perfect hash for %d samples", S);

    if (c->d_entries) { cudaFree(c->d_entries); c->d_entries = nullptr; }
    if (c->d_disp) { cudaFree(c->d_disp); c->d_disp = nullptr; }
    CU(cudaMalloc((void **)&c->d_entries, (size_t)n_slots * sizeof(uint4)));
    CU(cudaMalloc((void **)&c->d_disp, (size_t)n_buckets * sizeof(uint16_t)));
    CU(cudaMemcpy(c->d_entries, entries.data(), (size_t)n_slots * sizeof(uint4), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_disp, disp.data(), (size_t)n_buckets * sizeof(uint16_t), cudaMemcpyHostToDevice));
    c->table.entries = c->d_entries;
    c->table.disp = c->d_disp;
    c->table.n_buckets_mask = n_buckets - 1;
    c->table.n_slots_mask = n_slots - 1;
    c->table.seed = seed;
    for (auto &sl : c->slots) sl.dev.table = c->table;
    c->samples_set = true;
    return QUADE_OK;
}

#include "quade_api2.inc"
