/*
 * quade_oracle.c -- plain-C restatement of the reference hot path (see quade_oracle.h for status).
 *
 * It deliberately follows the reference's control flow -- four lock-step record cursors, one dict-like
 * lookup per pair, per-bin append buffers -- rather than the GPU design, so that the two implementations
 * fail differently.  Citations are into /root/reference.
 */
#include "quade_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* Python 2 str whitespace: space \t \n \v \f \r  (used by rstrip()/split(), FastqReader.py:42-44) */
static inline int is_ws(uint8_t c) { return c == 0x20 || (c >= 0x09 && c <= 0x0d); }

typedef struct {
    const uint8_t *b;
    uint64_t n;
    int last;
    uint64_t pos;     /* next unread byte */
    uint64_t skipped; /* invalid records passed so far */
    int blank_seen;   /* a record with no name token was passed (pending fatal error) */
} cursor;

typedef struct {
    uint64_t name_off, seq_off, qual_off;
    uint32_t name_len, seq_len;
} record;

/* One `next(fastq)` x4 + FastqSeq construction: FastqReader.py:37-44, FastqSeq.py:43-55.
 * returns 1 record parsed (valid flag in *valid), 0 exhausted. */
static int next_record(cursor *c, record *r, int *valid)
{
    uint64_t ls[4], le[4], p = c->pos;
    for (int i = 0; i < 4; i++) {
        if (p >= c->n) return 0;                       /* StopIteration, FastqReader.py:57 */
        const uint8_t *q = memchr(c->b + p, '\n', c->n - p);
        ls[i] = p;
        if (q) {
            le[i] = (uint64_t)(q - c->b);
            p = le[i] + 1;
        } else {
            if (!c->last) return 0;                    /* incomplete line: wait for the next batch */
            le[i] = c->n;                              /* final line without '\n' still is a line */
            p = c->n;
        }
    }
    /* name = line0.rstrip()[1:].split()[0]   FastqReader.py:42 */
    uint64_t e = le[0];
    while (e > ls[0] && is_ws(c->b[e - 1])) e--;
    uint64_t s = ls[0] + 1;                            /* [1:] drops the first byte whatever it is */
    while (s < e && is_ws(c->b[s])) s++;
    uint64_t t = s;
    while (t < e && !is_ws(c->b[t])) t++;
    r->name_off = s;
    r->name_len = (s < e) ? (uint32_t)(t - s) : 0;
    if (r->name_len == 0) c->blank_seen = 1;           /* IndexError in the reference */
    /* seq = line1.rstrip(); qual = line3.rstrip()   FastqReader.py:43-44 */
    e = le[1];
    while (e > ls[1] && is_ws(c->b[e - 1])) e--;
    r->seq_off = ls[1];
    r->seq_len = (uint32_t)(e - ls[1]);
    e = le[3];
    while (e > ls[3] && is_ws(c->b[e - 1])) e--;
    r->qual_off = ls[3];
    *valid = ((uint64_t)r->seq_len == e - ls[3]);      /* assert len(seq)==len(qual)  FastqSeq.py:55 */
    c->pos = p;
    return 1;
}

/* generator step: skip invalid records of THIS stream only (FastqReader.py:48-50).
 * returns 1 = valid record, 0 = exhausted, -1 = a parsed record had no name token: the reference raises
 * IndexError at FastqReader.py:42, i.e. BEFORE the length assertion and whatever the other streams hold. */
static int next_valid(cursor *c, record *r)
{
    int valid;
    while (next_record(c, r, &valid)) {
        if (c->blank_seen) return -1;
        if (valid) return 1;
        c->skipped++;
    }
    return 0;
}

/* passes the complete invalid records in front of the next valid one; stops before a valid record, before a record
 * without a name token, and where the buffered complete records end */
static void skip_invalid(cursor *c)
{
    for (;;) {
        cursor t = *c;
        record r;
        int valid;
        if (!next_record(&t, &r, &valid) || t.blank_seen || valid) return;
        t.skipped++;
        *c = t;
    }
}

/* ---- sample registry: INDEX_TO_SAMPLE dict (Sample.py:37,151) as an open-addressing table ---- */
typedef struct {
    int32_t *slot; /* sample id or -1 */
    uint32_t mask;
    const qo_params *p;
} registry;

static uint64_t fnv1a(const uint8_t *s, uint32_t n)
{
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < n; i++) { h ^= s[i]; h *= 1099511628211ull; }
    return h;
}

static int registry_build(registry *g, const qo_params *p)
{
    uint32_t cap = 16;
    while (cap < (uint32_t)p->n_samples * 4u) cap <<= 1;
    g->slot = (int32_t *)malloc(sizeof(int32_t) * cap);
    if (!g->slot) return QO_ERR_NOMEM;
    for (uint32_t i = 0; i < cap; i++) g->slot[i] = -1;
    g->mask = cap - 1;
    g->p = p;
    for (int32_t s = 0; s < p->n_samples; s++) {
        const uint8_t *k = p->keys + p->key_off[s];
        uint32_t len = p->key_off[s + 1] - p->key_off[s];
        if (len == 0) return QO_ERR_BAD_SAMPLE;        /* reject the empty index */
        for (uint32_t i = 0; i < len; i++)             /* Sample._is_dna, Sample.py:171-175 */
            if (!(k[i] == 'A' || k[i] == 'T' || k[i] == 'C' || k[i] == 'G' || k[i] == 'N')) return QO_ERR_BAD_SAMPLE;
        uint32_t h = (uint32_t)fnv1a(k, len) & g->mask;
        while (g->slot[h] >= 0) {
            int32_t o = g->slot[h];
            uint32_t ol = p->key_off[o + 1] - p->key_off[o];
            if (ol == len && memcmp(p->keys + p->key_off[o], k, len) == 0) return QO_ERR_BAD_SAMPLE; /* Sample.py:140 */
            h = (h + 1) & g->mask;
        }
        g->slot[h] = s;
    }
    return QO_OK;
}

static int32_t registry_find(const registry *g, const uint8_t *k, uint32_t len)
{
    uint32_t h = (uint32_t)fnv1a(k, len) & g->mask;
    while (g->slot[h] >= 0) {
        int32_t o = g->slot[h];
        uint32_t ol = g->p->key_off[o + 1] - g->p->key_off[o];
        if (ol == len && memcmp(g->p->keys + g->p->key_off[o], k, len) == 0) return o;
        h = (h + 1) & g->mask;
    }
    return -1;
}

/* ---- per-bin append buffers: FastqWriter.R1_buffer / R2_buffer (FastqWriter.py:33-34,68-69) ---- */
typedef struct { uint8_t *d; uint64_t n, cap; } bytebuf;

static int bb_reserve(bytebuf *b, uint64_t extra)
{
    if (b->n + extra <= b->cap) return 0;
    uint64_t nc = b->cap ? b->cap * 2 : 4096;
    while (nc < b->n + extra) nc *= 2;
    uint8_t *nd = (uint8_t *)realloc(b->d, nc);
    if (!nd) return -1;
    b->d = nd; b->cap = nc;
    return 0;
}

/* python slice clamp: s[a:b] on a string of length L (FastqSeq.__getitem__, FastqSeq.py:91-93) */
static inline void clamp_span(const int32_t span[2], uint32_t L, uint32_t *a, uint32_t *b)
{
    uint32_t s = (uint32_t)span[0], e = (uint32_t)span[1];
    *a = s < L ? s : L;
    *b = e < L ? e : L;
    if (*b < *a) *b = *a;
}

/* FastqWriter.__call__ (FastqWriter.py:61-69) + FastqSeq.fastqstr (FastqSeq.py:76-82) for one read */
static int emit(bytebuf *o, const uint8_t *rb, const record *r, const uint8_t *idx, uint32_t idx_len,
                const uint8_t *mol, uint32_t mol_len)
{
    uint64_t need = 1 + (uint64_t)r->name_len + 1 + idx_len + (mol_len ? 1 + mol_len : 0) + 1 +
                    (uint64_t)r->seq_len + 3 + r->seq_len + 1;
    if (bb_reserve(o, need)) return -1;
    uint8_t *w = o->d + o->n;
    *w++ = '@';
    memcpy(w, rb + r->name_off, r->name_len); w += r->name_len;
    *w++ = ':';
    memcpy(w, idx, idx_len); w += idx_len;             /* raw case, FastqWriter.py:62-66 */
    if (mol_len) { *w++ = ':'; memcpy(w, mol, mol_len); w += mol_len; }
    *w++ = '\n';
    memcpy(w, rb + r->seq_off, r->seq_len); w += r->seq_len;
    *w++ = '\n'; *w++ = '+'; *w++ = '\n';              /* separator normalised, descr dropped */
    memcpy(w, rb + r->qual_off, r->seq_len); w += r->seq_len; /* chr(ord(x)-33+33) == x */
    *w++ = '\n';
    o->n += need;
    return 0;
}

int qo_demux_batch(const qo_params *p, const uint8_t *const stream[4], const uint64_t nbytes[4], int last,
                   qo_result *res)
{
    memset(res, 0, sizeof(*res));
    const int ns = p->dual ? 4 : 3;
    const int nbins = 1 + 2 * p->n_samples;
    res->nbins = nbins;
    registry reg; memset(&reg, 0, sizeof(reg));
    bytebuf *bufs = (bytebuf *)calloc((size_t)nbins * 2, sizeof(bytebuf));
    res->counters = (int64_t *)calloc((size_t)nbins, sizeof(int64_t));
    res->bin_off[0] = (uint64_t *)calloc((size_t)nbins + 1, sizeof(uint64_t));
    res->bin_off[1] = (uint64_t *)calloc((size_t)nbins + 1, sizeof(uint64_t));
    int err = QO_OK;
    if (!bufs || !res->counters || !res->bin_off[0] || !res->bin_off[1]) { err = QO_ERR_NOMEM; goto done; }
    if ((err = registry_build(&reg, p)) != QO_OK) goto done;

    cursor cur[4]; memset(cur, 0, sizeof(cur));
    for (int s = 0; s < ns; s++) { cur[s].b = stream[s]; cur[s].n = nbytes[s]; cur[s].last = (last & 1) || ((last >> (8 + s)) & 1); }

    uint8_t keybuf[4096], molbuf[4096];
    for (;;) {                                         /* while True:  Quade.py:210 / :240 */
        record rec[4];
        cursor save[4];
        memcpy(save, cur, sizeof(cur));
        int got = 1, sfail = -1;
        for (int s = 0; s < ns && got > 0; s++) { got = next_valid(&cur[s], &rec[s]); sfail = s; }   /* Quade.py:211-214 */
        if (got < 0) { err = QO_ERR_BLANK_HEADER; goto done; }
        if (!got) {
            if (cur[sfail].last) { res->chunk_end = 1; break; }   /* StopIteration: end of chunk, Quade.py:223 */
            /* Batch semantics (the reference has no batches): the BUFFER of file sfail ran out before its end-of-file.
             * The unfinished iteration is given back: the caller carries every tail into the next batch.  What stays
             * consumed are the invalid records in front of each file's next valid record -- the reference skips them
             * in their own file whatever the other files hold (FastqReader.py:48-50) -- so that a buffer holding
             * nothing but such records makes progress; a blank header the reference may never reach is not passed. */
            for (int s = 0; s < ns; s++) {
                cur[s] = save[s];
                skip_invalid(&cur[s]);
                res->consumed[s] = cur[s].pos;
                res->n_skipped[s] = cur[s].skipped;
            }
            break;
        }
        for (int s = 0; s < ns; s++) { res->consumed[s] = cur[s].pos; res->n_skipped[s] = cur[s].skipped; }
        res->n_pairs++;                                /* Sample.TOTAL += 1, Sample.py:62 */

        /* index / molecular = I1[a:b] + I2[a:b]   Quade.py:217-218 (dual) / :246-247 (single) */
        uint32_t a, b, idx_len = 0, mol_len = 0;
        int minq = 1 << 30;
        const int ispan[2] = { QO_SPAN_IDX1, QO_SPAN_IDX2 }, mspan[2] = { QO_SPAN_MOL1, QO_SPAN_MOL2 };
        for (int k = 0; k < (p->dual ? 2 : 1); k++) {
            const uint8_t *ib = stream[QO_I1 + k];
            const record *ir = &rec[QO_I1 + k];
            clamp_span(p->span[ispan[k]], ir->seq_len, &a, &b);
            for (uint32_t i = a; i < b && idx_len < sizeof(keybuf); i++) {
                keybuf[idx_len++] = ib[ir->seq_off + i];
                int q = (int)ib[ir->qual_off + i] - 33;   /* FastqSeq.py:44 */
                if (q < minq) minq = q;
            }
            clamp_span(p->span[mspan[k]], ir->seq_len, &a, &b);
            for (uint32_t i = a; i < b && mol_len < sizeof(molbuf); i++) molbuf[mol_len++] = ib[ir->seq_off + i];
        }
        /* index.seq.upper() in INDEX_TO_SAMPLE   Sample.py:65 */
        uint8_t up[4096];
        for (uint32_t i = 0; i < idx_len; i++) up[i] = (keybuf[i] >= 'a' && keybuf[i] <= 'z') ? keybuf[i] - 32 : keybuf[i];
        int32_t sid = registry_find(&reg, up, idx_len);
        int bin, wr;
        if (sid >= 0) {
            if (minq >= p->min_qual) { bin = 1 + 2 * sid; wr = p->write_pass; }   /* Sample.py:70-75 */
            else                     { bin = 2 + 2 * sid; wr = p->write_fail; }   /* Sample.py:78-83 */
        } else { bin = 0; wr = p->write_undetermined; }                           /* Sample.py:86-91 */
        res->counters[bin]++;
        if (wr) {
            if (emit(&bufs[bin * 2 + 0], stream[QO_R1], &rec[QO_R1], keybuf, idx_len, molbuf, mol_len) ||
                emit(&bufs[bin * 2 + 1], stream[QO_R2], &rec[QO_R2], keybuf, idx_len, molbuf, mol_len)) {
                err = QO_ERR_NOMEM; goto done;
            }
        }
    }
    /* lay the per-bin buffers out as one arena per read, bins in order */
    for (int k = 0; k < 2; k++) {
        uint64_t tot = 0;
        for (int bnum = 0; bnum < nbins; bnum++) { res->bin_off[k][bnum] = tot; tot += bufs[bnum * 2 + k].n; }
        res->bin_off[k][nbins] = tot;
        res->out[k] = (uint8_t *)malloc(tot ? tot : 1);
        if (!res->out[k]) { err = QO_ERR_NOMEM; goto done; }
        for (int bnum = 0; bnum < nbins; bnum++)
            if (bufs[bnum * 2 + k].n) memcpy(res->out[k] + res->bin_off[k][bnum], bufs[bnum * 2 + k].d, bufs[bnum * 2 + k].n);
    }
done:
    if (bufs) { for (int i = 0; i < nbins * 2; i++) free(bufs[i].d); free(bufs); }
    free(reg.slot);
    res->err = err;
    return err;
}

void qo_result_free(qo_result *res)
{
    free(res->out[0]); free(res->out[1]);
    free(res->bin_off[0]); free(res->bin_off[1]);
    free(res->counters);
    memset(res, 0, sizeof(*res));
}

/* ---- multi-threaded chunk runner for the CPU baseline ---- */
typedef struct {
    const qo_params *p;
    int n_chunks;
    const uint8_t *const *streams;
    const uint64_t *nbytes;
    int next;
    pthread_mutex_t mu;
    int64_t pairs;
    uint64_t out_bytes;
    int err;
} mt_job;

static void *mt_worker(void *arg)
{
    mt_job *j = (mt_job *)arg;
    for (;;) {
        pthread_mutex_lock(&j->mu);
        int c = j->next++;
        pthread_mutex_unlock(&j->mu);
        if (c >= j->n_chunks) break;
        qo_result r;
        int e = qo_demux_batch(j->p, j->streams + 4 * c, j->nbytes + 4 * c, 1, &r);
        pthread_mutex_lock(&j->mu);
        if (e) j->err = e;
        j->pairs += (int64_t)r.n_pairs;
        if (!e) j->out_bytes += r.bin_off[0][r.nbins] + r.bin_off[1][r.nbins];
        pthread_mutex_unlock(&j->mu);
        qo_result_free(&r);
    }
    return NULL;
}

int64_t qo_demux_chunks_mt(const qo_params *p, int n_chunks, const uint8_t *const *streams, const uint64_t *nbytes,
                           int n_threads, double *seconds_out, uint64_t *out_bytes_total)
{
    mt_job j; memset(&j, 0, sizeof(j));
    j.p = p; j.n_chunks = n_chunks; j.streams = streams; j.nbytes = nbytes;
    pthread_mutex_init(&j.mu, NULL);
    if (n_threads < 1) n_threads = 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < n_threads; i++) pthread_create(&th[i], NULL, mt_worker, &j);
    for (int i = 0; i < n_threads; i++) pthread_join(th[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th);
    pthread_mutex_destroy(&j.mu);
    if (seconds_out) *seconds_out = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    if (out_bytes_total) *out_bytes_total = j.out_bytes;
    return j.err ? (int64_t)j.err : j.pairs;
}
