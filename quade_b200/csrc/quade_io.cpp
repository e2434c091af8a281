// quade_io.cpp -- libquade_io.so (include/quade_io.h): parallel gzip-member writer for the per-bin output files and
// the inflating reader of gzip input files (src/pyFastq/FastqReader.py:30).
// Host C++ only (zlib, threads); no CUDA.  Replaces the file side of src/FastqWriter.py:83-90.
#include "../../include/quade_io.h"

#include <zlib.h>

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

struct Job {
    uint32_t file;          // 2 * bin + read
    uint64_t seq;           // position of this member among the members of its file
    const uint8_t *ptr;
    uint64_t len;
    uint64_t batch;
};

struct FileState {
    std::string path;
    uint64_t next_submit = 0;   // members handed out
    uint64_t next_write = 0;    // members on disk
};

} // namespace

struct qio_sink {
    int level = 1;
    std::vector<std::thread> workers;
    std::mutex mu;                       // queue, file order, batch accounting
    std::condition_variable cv_job;      // a job was queued, or stop
    std::condition_variable cv_order;    // a member reached its file: successors may write
    std::condition_variable cv_done;     // a batch finished
    std::deque<Job> queue;
    std::vector<FileState> files;
    std::deque<uint64_t> open_jobs;      // jobs left per batch; front = batch `first_open`
    uint64_t first_open = 1, next_batch = 1;
    bool stop = false;
    std::string error;
    int error_code = QIO_OK;
    std::atomic<uint64_t> bytes_in{0}, bytes_out{0};
};

namespace {

void worker_main(qio_sink *s)
{
    z_stream z;
    memset(&z, 0, sizeof(z));
    bool z_ok = deflateInit2(&z, s->level, Z_DEFLATED, 15 + 16 /* gzip member */, 8, Z_DEFAULT_STRATEGY) == Z_OK;
    std::vector<uint8_t> buf;
    for (;;) {
        Job j;
        {
            std::unique_lock<std::mutex> lk(s->mu);
            s->cv_job.wait(lk, [&] { return s->stop || !s->queue.empty(); });
            if (s->queue.empty()) break;
            j = s->queue.front();
            s->queue.pop_front();
        }
        std::string err;
        int code = QIO_OK;
        size_t produced = 0;
        if (!z_ok) {
            err = "deflateInit2 failed";
            code = QIO_ERR_ZLIB;
        } else {
            deflateReset(&z);
            // zlib counts in 32 bits per call: feed the range in pieces (a bin range is < 4 GiB: the arena is)
            size_t bound = 64;
            for (uint64_t o = 0; o < j.len; o += 1u << 30) bound += (size_t)deflateBound(&z, (uLong)std::min<uint64_t>(j.len - o, 1u << 30));
            buf.resize(bound);
            uint64_t off = 0;
            z.next_out = buf.data();
            z.avail_out = (uInt)std::min<size_t>(buf.size(), 0xffffffffu);
            int rc = Z_OK;
            do {
                const uint64_t piece = std::min<uint64_t>(j.len - off, 1u << 30);
                z.next_in = const_cast<Bytef *>(j.ptr + off);
                z.avail_in = (uInt)piece;
                off += piece;
                rc = deflate(&z, off == j.len ? Z_FINISH : Z_NO_FLUSH);
            } while (rc == Z_OK && off < j.len);
            if (rc != Z_STREAM_END) {
                err = "deflate failed";
                code = QIO_ERR_ZLIB;
            }
            produced = (size_t)(z.next_out - buf.data());
        }
        {
            // members of one file reach it in submission order (FastqWriter.py:68-69: records of a bin are in arrival order)
            std::unique_lock<std::mutex> lk(s->mu);
            s->cv_order.wait(lk, [&] { return s->files[j.file].next_write == j.seq; });
        }
        if (code == QIO_OK) {
            const std::string &path = s->files[j.file].path;      // set_paths waits for the jobs in flight: stable
            const int fd = open(path.c_str(), O_WRONLY | O_APPEND | O_CLOEXEC);
            if (fd < 0) {
                err = "open " + path + ": " + strerror(errno);
                code = QIO_ERR_IO;
            } else {
                size_t done = 0;
                while (done < produced) {
                    const ssize_t w = write(fd, buf.data() + done, produced - done);
                    if (w < 0) {
                        if (errno == EINTR) continue;
                        err = "write " + path + ": " + strerror(errno);
                        code = QIO_ERR_IO;
                        break;
                    }
                    done += (size_t)w;
                }
                close(fd);
            }
        }
        s->bytes_in += j.len;
        s->bytes_out += produced;
        {
            std::lock_guard<std::mutex> lk(s->mu);
            s->files[j.file].next_write = j.seq + 1;
            if (code != QIO_OK && s->error_code == QIO_OK) {
                s->error_code = code;
                s->error = err;
            }
            uint64_t &left = s->open_jobs[(size_t)(j.batch - s->first_open)];
            --left;
            while (!s->open_jobs.empty() && s->open_jobs.front() == 0) {
                s->open_jobs.pop_front();
                ++s->first_open;
            }
        }
        s->cv_order.notify_all();
        s->cv_done.notify_all();
    }
    if (z_ok) deflateEnd(&z);
}

int wait_locked(qio_sink *s, std::unique_lock<std::mutex> &lk, uint64_t ticket)
{
    s->cv_done.wait(lk, [&] { return s->first_open > ticket; });
    if (s->error_code != QIO_OK) return fail(s->error_code, "%s", s->error.c_str());
    return QIO_OK;
}

} // namespace

extern "C" const char *qio_last_error(void) { return g_err; }

extern "C" int qio_sink_create(int n_threads, int level, qio_sink **out)
{
    if (!out || level < 0 || level > 9) return fail(QIO_ERR_ARG, "out is NULL or level not in 0..9");
    if (n_threads <= 0) n_threads = (int)std::min(64u, std::max(1u, std::thread::hardware_concurrency()));
    qio_sink *s = new qio_sink();
    s->level = level;
    for (int i = 0; i < n_threads; ++i) s->workers.emplace_back(worker_main, s);
    *out = s;
    return QIO_OK;
}

extern "C" int qio_sink_destroy(qio_sink *s)
{
    if (!s) return QIO_OK;
    {
        std::unique_lock<std::mutex> lk(s->mu);
        s->cv_done.wait(lk, [&] { return s->open_jobs.empty(); });
        s->stop = true;
    }
    s->cv_job.notify_all();
    for (auto &t : s->workers) t.join();
    delete s;
    return QIO_OK;
}

extern "C" int qio_sink_set_paths(qio_sink *s, int n_bins, const char *const r1[], const char *const r2[])
{
    if (!s || n_bins < 0 || (n_bins && (!r1 || !r2))) return fail(QIO_ERR_ARG, "sink/paths is NULL");
    std::unique_lock<std::mutex> lk(s->mu);
    s->cv_done.wait(lk, [&] { return s->open_jobs.empty(); });
    s->files.assign((size_t)2 * n_bins, FileState());
    for (int b = 0; b < n_bins; ++b) {
        s->files[2 * b].path = r1[b];
        s->files[2 * b + 1].path = r2[b];
    }
    return QIO_OK;
}

extern "C" int qio_sink_append(qio_sink *s, const uint8_t *out1, const uint64_t *off1, const uint8_t *out2, const uint64_t *off2,
                               uint64_t *ticket)
{
    if (!s || !off1 || !off2 || !ticket) return fail(QIO_ERR_ARG, "sink/offsets/ticket is NULL");
    const size_t nb = s->files.size() / 2;
    size_t n_jobs = 0;
    {
        std::lock_guard<std::mutex> lk(s->mu);
        const uint64_t batch = s->next_batch++;
        s->open_jobs.push_back(0);
        uint64_t &left = s->open_jobs.back();
        const uint8_t *base[2] = {out1, out2};
        const uint64_t *off[2] = {off1, off2};
        for (size_t b = 0; b < nb; ++b)
            for (int k = 0; k < 2; ++k) {
                const uint64_t lo = off[k][b], hi = off[k][b + 1];
                if (hi <= lo) continue;
                FileState &f = s->files[2 * b + k];
                s->queue.push_back(Job{(uint32_t)(2 * b + k), f.next_submit++, base[k] + lo, hi - lo, batch});
                ++left;
                ++n_jobs;
            }
        *ticket = batch;
        if (left == 0) {                                    // an empty batch is done at once (if nothing older is open)
            while (!s->open_jobs.empty() && s->open_jobs.front() == 0) {
                s->open_jobs.pop_front();
                ++s->first_open;
            }
        }
    }
    if (n_jobs) s->cv_job.notify_all();
    else s->cv_done.notify_all();
    return QIO_OK;
}

extern "C" int qio_sink_wait(qio_sink *s, uint64_t ticket)
{
    if (!s) return fail(QIO_ERR_ARG, "sink is NULL");
    std::unique_lock<std::mutex> lk(s->mu);
    return wait_locked(s, lk, ticket);
}

extern "C" int qio_sink_stats(qio_sink *s, uint64_t *bytes_in, uint64_t *bytes_out)
{
    if (!s) return fail(QIO_ERR_ARG, "sink is NULL");
    if (bytes_in) *bytes_in = s->bytes_in.load();
    if (bytes_out) *bytes_out = s->bytes_out.load();
    return QIO_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// reader: one inflating thread per gzip file, a ring of decompressed blocks
// ---------------------------------------------------------------------------------------------------------------
struct qio_reader {
    int fd = -1;
    std::string path;
    size_t block_bytes = 0;
    std::thread worker;
    std::mutex mu;
    std::condition_variable cv_full, cv_free;      // a block was filled / a block was handed back
    std::deque<std::vector<uint8_t>> full;         // filled blocks, oldest first
    std::vector<std::vector<uint8_t>> spare;       // blocks to fill
    bool done = false, stop = false;               // producer finished (end of stream or error) / consumer closes
    int error_code = QIO_OK;
    std::string error;
    std::vector<uint8_t> cur;                      // block being drained by the consumer
    size_t cur_off = 0;
};

namespace {

void reader_main(qio_reader *r)
{
    z_stream z;
    memset(&z, 0, sizeof(z));
    std::string err;
    int code = QIO_OK;
    if (inflateInit2(&z, 15 + 32 /* gzip or zlib header, detected */) != Z_OK) {
        err = "inflateInit2 failed";
        code = QIO_ERR_ZLIB;
    }
    std::vector<uint8_t> in(1u << 20);
    bool in_member = false, any_member = false, eof_in = false;
    std::vector<uint8_t> blk;
    size_t filled = 0;
    auto take_spare = [&]() -> bool {               // false: the consumer is closing
        std::unique_lock<std::mutex> lk(r->mu);
        r->cv_free.wait(lk, [&] { return r->stop || !r->spare.empty(); });
        if (r->stop) return false;
        blk = std::move(r->spare.back());
        r->spare.pop_back();
        blk.resize(r->block_bytes);
        filled = 0;
        return true;
    };
    auto push_block = [&]() {
        blk.resize(filled);
        {
            std::lock_guard<std::mutex> lk(r->mu);
            r->full.push_back(std::move(blk));
        }
        r->cv_full.notify_one();
        blk = std::vector<uint8_t>();
        filled = 0;
    };
    bool have_blk = code == QIO_OK && take_spare();
    while (code == QIO_OK && have_blk) {
        if (z.avail_in == 0 && !eof_in) {
            ssize_t n;
            do n = read(r->fd, in.data(), in.size()); while (n < 0 && errno == EINTR);
            if (n < 0) {
                err = "read " + r->path + ": " + strerror(errno);
                code = QIO_ERR_IO;
                break;
            }
            if (n == 0) eof_in = true;
            z.next_in = in.data();
            z.avail_in = (uInt)n;
        }
        if (!in_member) {
            // between members: zero padding is skipped (Python's gzip does), anything else must be a new member
            while (z.avail_in && *z.next_in == 0 && any_member) { ++z.next_in; --z.avail_in; }
            if (z.avail_in == 0) {
                if (eof_in) break;                   // clean end of the stream
                continue;
            }
            if (any_member) inflateReset(&z);
            in_member = any_member = true;
        }
        if (z.avail_in == 0 && eof_in) {
            err = "unexpected end of file inside a gzip member: " + r->path;
            code = QIO_ERR_ZLIB;
            break;
        }
        z.next_out = blk.data() + filled;
        z.avail_out = (uInt)std::min<size_t>(r->block_bytes - filled, 0x40000000u);
        const uInt before = z.avail_out;
        const int rc = inflate(&z, Z_NO_FLUSH);
        filled += before - z.avail_out;
        if (rc == Z_STREAM_END) in_member = false;
        else if (rc != Z_OK && rc != Z_BUF_ERROR) {
            err = std::string("inflate failed (") + (z.msg ? z.msg : "corrupt data") + "): " + r->path;
            code = QIO_ERR_ZLIB;
            break;
        }
        if (filled == r->block_bytes) {
            push_block();
            have_blk = take_spare();
        }
    }
    if (have_blk && filled) push_block();            // the bytes in front of an error are still good
    inflateEnd(&z);
    {
        std::lock_guard<std::mutex> lk(r->mu);
        r->done = true;
        if (code != QIO_OK) {
            r->error_code = code;
            r->error = err;
        }
    }
    r->cv_full.notify_all();
}

} // namespace

extern "C" int qio_reader_open(const char *path, uint64_t block_bytes, int depth, qio_reader **out)
{
    if (!path || !out || block_bytes < 4096 || block_bytes > (1ull << 30) || depth < 1 || depth > 4096)
        return fail(QIO_ERR_ARG, "path/out is NULL, or block_bytes not in 4 KiB..1 GiB, or depth not in 1..4096");
    const int fd = open(path, O_RDONLY | O_CLOEXEC);
    if (fd < 0) return fail(QIO_ERR_IO, "open %s: %s", path, strerror(errno));
    qio_reader *r = new qio_reader();
    r->fd = fd;
    r->path = path;
    r->block_bytes = (size_t)block_bytes;
    r->spare.resize((size_t)depth + 1);              // blocks are sized when they are first filled
    r->worker = std::thread(reader_main, r);
    *out = r;
    return QIO_OK;
}

extern "C" int qio_reader_read(qio_reader *r, uint8_t *dst, uint64_t cap, uint64_t *got)
{
    if (!r || !got || (cap && !dst)) return fail(QIO_ERR_ARG, "reader/dst/got is NULL");
    uint64_t done = 0;
    while (done < cap) {
        if (r->cur_off == r->cur.size()) {
            std::unique_lock<std::mutex> lk(r->mu);
            if (r->cur.capacity()) {                                 // hand the drained block back (its size stays: no refill)
                r->spare.push_back(std::move(r->cur));
                r->cur = std::vector<uint8_t>();
                r->cv_free.notify_one();
            }
            r->cur_off = 0;
            if (done) {                                              // never wait with bytes in hand
                if (r->full.empty()) break;
            } else {
                r->cv_full.wait(lk, [&] { return r->done || !r->full.empty(); });
            }
            if (r->full.empty()) {                                   // producer finished
                if (r->error_code != QIO_OK) {
                    *got = 0;
                    return fail(r->error_code, "%s", r->error.c_str());
                }
                break;
            }
            r->cur = std::move(r->full.front());
            r->full.pop_front();
        }
        const size_t n = (size_t)std::min<uint64_t>(cap - done, r->cur.size() - r->cur_off);
        memcpy(dst + done, r->cur.data() + r->cur_off, n);
        r->cur_off += n;
        done += n;
    }
    *got = done;
    return QIO_OK;
}

extern "C" int qio_reader_close(qio_reader *r)
{
    if (!r) return QIO_OK;
    {
        std::lock_guard<std::mutex> lk(r->mu);
        r->stop = true;
    }
    r->cv_free.notify_all();
    if (r->worker.joinable()) r->worker.join();
    if (r->fd >= 0) close(r->fd);
    delete r;
    return QIO_OK;
}
