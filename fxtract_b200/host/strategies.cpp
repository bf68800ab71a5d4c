// strategies.cpp -- the three literal strategies of the reference (main.cpp:306-349 hash_search,
// :351-452 multipattern_search, :596-652 simple_string_search) with every matcher on the GPU.
//
// Two front-ends feed the same device matchers:
//   raw path    -- canonical uncompressed single-end FASTQ: raw bytes go to the GPU, which finds
//                  record boundaries, validates, matches and returns the selected byte runs; the
//                  host writes those runs from its own pinned copy.  A batch the GPU predicate
//                  rejects is re-routed, from its first byte, to the
//   fields path -- host reader with kseq semantics -> target fields packed into the pinned
//                  buffer -> per-field flags (+ first-hit pattern ids) -> per-strategy pair rules,
//                  printing and routing on the host exactly as the reference does.
// The POSIX-regex strategies (main.cpp:462-514) are outside the graft and stay on the CPU.
#include "fxhost.h"
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cerrno>
#include <cstdlib>
#include <mutex>
#include <set>
#include <thread>
#include <cstring>
#include <fcntl.h>
#include <memory>
#include <regex.h>
#include <sys/stat.h>
#include <sys/uio.h>
#include <unistd.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#ifndef REG_BASIC
#define REG_BASIC 0
#endif

namespace fxh {

namespace {

enum Strategy { S_LITERAL, S_MULTI, S_EXACT };

struct Stats {
    uint64_t raw_bytes = 0, raw_batches = 0, field_bytes = 0, field_batches = 0, records = 0, rerouted = 0;
    double t_ingest = 0, t_gpu = 0, t_write = 0;      // raw path: main thread blocked on the readers / on the devices / writing
    double t_submit = 0, t_setup = 0;                 // ... inside the submit calls (copy + kernel enqueue) / allocating the staging windows
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Selected byte runs of a raw batch go out with writev straight from the pinned copy of the batch (stdio would
// copy every byte once more into its buffer: with -v nearly the whole input is output).
void write_runs(const uint8_t *host, const fxg_run *runs, uint64_t n)
{
    if (!n) return;
    fflush(stdout);                                     // keep the order with anything printed through stdio
    constexpr int kIov = 512;
    struct iovec iov[kIov];
    uint64_t k = 0;
    while (k < n) {
        int cnt = 0;
        size_t want = 0;
        for (; cnt < kIov && k < n; ++k) {
            if (!runs[k].len) continue;
            iov[cnt].iov_base = const_cast<uint8_t *>(host + runs[k].off);
            iov[cnt].iov_len = (size_t)runs[k].len;
            want += iov[cnt].iov_len;
            ++cnt;
        }
        int first = 0;
        while (want) {
            ssize_t got = writev(STDOUT_FILENO, iov + first, cnt - first);
            if (got < 0) { if (errno == EINTR) continue; perror("fxtract: write"); exit(1); }
            want -= (size_t)got;
            while (got > 0 && first < cnt) {              // short write: advance through the vector
                if ((size_t)got >= iov[first].iov_len) { got -= (ssize_t)iov[first].iov_len; ++first; }
                else { iov[first].iov_base = (uint8_t *)iov[first].iov_base + got; iov[first].iov_len -= (size_t)got; got = 0; }
            }
        }
    }
}

// main.cpp:291-304 prepare_data; the multi-pattern twin (main.cpp:389-405, :410-425) differs for -Q
void target_slice(const CliOptions &o, Strategy st, const PairBatch &b, const RecView &r, bool second,
                  const char *&p, uint32_t &n)
{
    Field f;
    bool none = false;
    if (o.H) f = F_NAME;
    else if (o.Q) {
        if (st == S_MULTI) {
            if (r.is_fasta()) none = true;            // reference leaves `data` untouched (UB / stale): nothing to match
            f = second ? F_SEQ : F_QUAL;              // main.cpp:416 scans read 2's *sequence* under -Q
        } else f = r.is_fasta() ? F_SEQ : F_QUAL;
    } else if (o.C) f = F_COMMENT;
    else f = F_SEQ;
    p = b.arena.data() + r.off[f];
    n = none ? 0 : r.len[f];
    if (st == S_MULTI && o.Q && !none) n = r.len[F_SEQ];     // data.len = pair.X.length()
    if (st == S_LITERAL) n = (uint32_t)strnlen(p, n);        // strstr works on c_str()
}

int fxg_target(const CliOptions &o)
{
    return o.H ? FXG_TARGET_NAME : (o.Q ? FXG_TARGET_QUAL : (o.C ? FXG_TARGET_COMMENT : FXG_TARGET_SEQ));
}

// on_match (main.cpp:258-271): 1 = "-l: caller prints the file name and stops"
int on_match(Job &job, FILE *out, const PairBatch &b, size_t i)
{
    if (job.opt.l) return 1;
    if (job.opt.c) { ++job.matched_records; return 0; }
    if (out) print_pair(out, b, i);
    return 0;
}

// Applies one strategy's pair rules to a batch whose per-field flags are known.
// Returns true when the input is finished early (-l).
bool apply_rules(Job &job, Strategy st, const PairBatch &b, bool paired, const uint8_t *flags, const int32_t *pid,
                 const std::string &file1, const std::string &file2)
{
    const CliOptions &o = job.opt;
    const size_t stride = paired ? 2 : 1;
    for (size_t i = 0; i < b.size(); ++i) {
        const bool m1 = flags[i * stride] != 0;
        const bool has2 = paired && !b.second[i].empty();
        const bool m2 = has2 && flags[i * stride + 1] != 0;
        if (st == S_LITERAL) {                                   // main.cpp:599-648
            if (m1) {
                if (!o.v && on_match(job, stdout, b, i) == 1) { printf("%s\n", file1.c_str()); return true; }
            } else if (o.v) on_match(job, stdout, b, i);
            else if (m2 && on_match(job, stdout, b, i) == 1) { printf("%s\n", file1.c_str()); return true; }   // :636 prints file 1
        } else if (st == S_EXACT) {                              // main.cpp:311-346
            if (m1) {
                FILE *out = job.router.find(pid[i * stride]);    // lookup side effects happen even under -v / -c
                if (!o.v && on_match(job, out, b, i) == 1) { printf("%s\n", file1.c_str()); return true; }
            } else if (has2) {
                if (m2) {
                    FILE *out = job.router.find(pid[i * stride + 1]);
                    if (on_match(job, out, b, i) == 1) { printf("%s\n", file2.c_str()); return true; }    // -v ignored here (:331-337)
                } else if (o.v) on_match(job, stdout, b, i);
            } else if (o.v) on_match(job, stdout, b, i);
        } else {                                                 // main.cpp:388-446 with on_match2 (:273-289)
            auto hit = [&](int id) {
                if (o.v || o.l) return;
                if (o.c) { ++job.matched_records; return; }
                FILE *out = job.router.find(id);
                if (out) print_pair(out, b, i);
            };
            if (m1) {
                hit(pid[i * stride]);
                if (o.l) { printf("%s\n", file1.c_str()); return true; }
            } else if (has2) {
                if (m2) hit(pid[i * stride + 1]);
                if (!m2 && o.v) on_match(job, stdout, b, i);
                else if (o.l) { printf("%s\n", file2.c_str()); return true; }     // also taken for a miss without -v (:427-435)
            } else if (o.v) on_match(job, stdout, b, i);
        }
    }
    return false;
}

// ------------------------------------------------------------------ fields path
bool run_fields(Job &job, Strategy st, PairSource &src, bool paired, DevicePool &pool, Stats &stats)
{
    const size_t max_pairs = 1u << 20;
    std::deque<std::unique_ptr<PairBatch>> pending;
    std::vector<uint32_t> off;
    auto finish_oldest = [&]() -> bool {
        fxg_result r;
        pool.collect(&r);
        std::unique_ptr<PairBatch> b = std::move(pending.front());
        pending.pop_front();
        return apply_rules(job, st, *b, paired, r.flags, r.pattern_id, src.name1(), src.name2());
    };
    for (;;) {
        std::unique_ptr<PairBatch> b(new PairBatch());
        src.read_batch(*b, max_pairs, pool.batch_bytes() / 2);
        const bool ended = b->ended;
        if (b->size()) {
            if (b->arena.size() + 64 > pool.batch_bytes()) {        // one oversized record: drain, then enlarge the staging buffers
                while (!pending.empty()) if (finish_oldest()) return true;
                size_t want = pool.batch_bytes();
                while (want < b->arena.size() + 64) want <<= 1;
                pool.grow(want);
            }
            if (pool.inflight() == pool.capacity() && finish_oldest()) return true;
            void *pin; size_t cap;
            pool.acquire(&pin, &cap);
            char *dst = (char *)pin;
            size_t used = 0;
            off.clear();
            off.push_back(0);
            for (size_t i = 0; i < b->size(); ++i) {
                const char *p; uint32_t n;
                target_slice(job.opt, st, *b, b->first[i], false, p, n);
                memcpy(dst + used, p, n); used += n; off.push_back((uint32_t)used);
                if (paired) {
                    if (b->second[i].empty()) n = 0;
                    else target_slice(job.opt, st, *b, b->second[i], true, p, n);
                    memcpy(dst + used, p, n); used += n; off.push_back((uint32_t)used);
                }
            }
            pool.submit_fields(used, off.data(), (uint32_t)off.size() - 1);
            stats.field_bytes += used; stats.field_batches++; stats.records += b->size();
            pending.push_back(std::move(b));
        }
        if (ended) break;
    }
    while (!pending.empty()) if (finish_oldest()) return true;
    return false;
}

// ------------------------------------------------------------------ raw path
// '\n' count of a slice (mates on the raw path: the host cuts both streams at the same record count, and canonical
// FASTQ -- which the GPU predicate verifies -- has four lines per record)
uint64_t count_newlines(const uint8_t *p, size_t n)
{
    uint64_t cnt = 0;
    size_t i = 0;
#if defined(__SSE2__)
    const __m128i nl = _mm_set1_epi8('\n');
    for (; i + 64 <= n; i += 64) {
        const unsigned m0 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i)), nl));
        const unsigned m1 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i + 16)), nl));
        const unsigned m2 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i + 32)), nl));
        const unsigned m3 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i + 48)), nl));
        cnt += (uint64_t)__builtin_popcountll((unsigned long long)m0 | ((unsigned long long)m1 << 16) |
                                              ((unsigned long long)m2 << 32) | ((unsigned long long)m3 << 48));
    }
#endif
    for (; i < n; ++i) cnt += p[i] == '\n';
    return cnt;
}

// offset just behind the `skip`-th '\n' counted backwards from the end of buf[0, n) after dropping what follows the
// last one: with L '\n' in the buffer, skip = L - K gives the offset behind the K-th.  (K >= 1.)
size_t offset_behind_newline_from_end(const uint8_t *buf, size_t n, uint64_t skip)
{
    const uint8_t *e = (const uint8_t *)memrchr(buf, '\n', n);
    while (e && skip--) e = (e > buf) ? (const uint8_t *)memrchr(buf, '\n', (size_t)(e - buf)) : nullptr;
    return e ? (size_t)(e - buf) + 1 : 0;
}

// Host ingest of the raw path (hard part 5): the file is read in fixed windows by a team of
// reader threads (pread of 4 MiB slices) straight into pinned buffers, several windows ahead of the GPU.  A window
// becomes a batch once the unconsumed tail of its predecessor (a partial record, copied into the pad in front of
// the window) is known and the last record boundary inside it is found; batches go to the devices round-robin and
// come back in order, and the selected byte runs are written from the pinned copy.
class RawIngest {
public:
    static constexpr size_t kSlice = 4u << 20;
    // `pad` = room in front of every window for the carried partial record
    // `nwin` > 0: that many windows (the ones past the end of the file are empty) -- mates from two files advance
    // through both inputs in step
    RawIngest(int fd, uint64_t file_size, size_t window, size_t pad, int nbuf, int nthreads, bool count_nl = false, uint64_t nwin = 0)
        : fd_(fd), size_(file_size), window_(window), pad_(pad), nbuf_(nbuf), count_nl_(count_nl)
    {
        nwin_ = nwin ? nwin : (size_ + window_ - 1) / window_;
        bufs_.resize(nbuf_);
        for (Buf &b : bufs_) {
            void *p = nullptr;
            if (fxg_host_alloc(&p, pad_ + window_ + 64) != FXG_OK) { fprintf(stderr, "fxtract: cannot allocate pinned staging memory: %s\n", fxg_last_error(nullptr)); exit(1); }
            b.mem = (uint8_t *)p;
        }
        done_.assign(nwin_, 0);
        nl_.assign(nwin_, 0);
        for (int t = 0; t < nthreads; ++t) threads_.emplace_back([this] { work(); });
    }
    ~RawIngest()
    {
        { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
        cv_.notify_all();
        for (std::thread &t : threads_) t.join();
        for (Buf &b : bufs_) fxg_host_free(b.mem);
    }
    uint64_t windows() const { return nwin_; }
    uint64_t window_lines(uint64_t k) const { return nl_[k]; }       // '\n' count of window k (count_nl; after wait_window)
    // blocks until window k is in memory; returns its first byte (`pad` bytes lie in front of it) and length
    uint8_t *wait_window(uint64_t k, size_t *len)
    {
        std::unique_lock<std::mutex> l(mu_);
        const size_t want = slices_in(k);
        cv_.wait(l, [&] { return done_[k] == want || failed_; });
        if (failed_) { perror("fxtract: read"); exit(1); }
        *len = win_len(k);
        return bufs_[k % nbuf_].mem + pad_;
    }
    // window k's buffer may be overwritten (by window k + nbuf)
    void release_window(uint64_t k)
    {
        { std::lock_guard<std::mutex> l(mu_); released_.insert(k); while (released_.count(low_)) { released_.erase(low_); ++low_; } }
        cv_.notify_all();
    }
private:
    struct Buf { uint8_t *mem = nullptr; };
    size_t win_len(uint64_t k) const { const uint64_t o = k * window_; return o >= size_ ? 0 : (size_t)std::min<uint64_t>(window_, size_ - o); }
    size_t slices_in(uint64_t k) const { return (win_len(k) + kSlice - 1) / kSlice; }
    void work()
    {
        for (;;) {
            uint64_t k; size_t s;
            {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] {
                    while (next_win_ < nwin_ && slices_in(next_win_) == 0) ++next_win_;     // empty windows need no reader
                    return stop_ || failed_ || (next_win_ < nwin_ && next_win_ < low_ + (uint64_t)nbuf_);
                });
                if (stop_ || failed_) return;
                k = next_win_; s = next_slice_;
                if (++next_slice_ >= slices_in(k)) { next_slice_ = 0; ++next_win_; }
            }
            uint8_t *dst = bufs_[k % nbuf_].mem + pad_ + s * kSlice;
            const uint64_t off = k * window_ + (uint64_t)s * kSlice;
            size_t want = std::min(kSlice, win_len(k) - s * kSlice), have = 0;
            bool bad = false;
            while (have < want) {
                ssize_t got = pread(fd_, dst + have, want - have, (off_t)(off + have));
                if (got <= 0) { bad = true; break; }          // the file shrank under us, or an I/O error
                have += (size_t)got;
            }
            const uint64_t lines = (count_nl_ && !bad) ? count_newlines(dst, want) : 0;
            { std::lock_guard<std::mutex> l(mu_); if (bad) failed_ = true; nl_[k] += lines; ++done_[k]; }
            cv_.notify_all();
        }
    }
    int fd_; uint64_t size_; size_t window_, pad_; int nbuf_; bool count_nl_;
    uint64_t nwin_ = 0;
    std::vector<uint64_t> nl_;
    std::vector<Buf> bufs_;
    std::vector<size_t> done_;
    std::vector<std::thread> threads_;
    std::mutex mu_; std::condition_variable cv_;
    uint64_t next_win_ = 0, low_ = 0; size_t next_slice_ = 0;     // low_: every window below it has been released
    std::set<uint64_t> released_;
    bool stop_ = false, failed_ = false;
};

struct RawPending { const uint8_t *host; uint64_t file_off; size_t nbytes; uint64_t window; };

// Which form of the raw path an input may take.  Per-pattern output files need the pattern that selected a
// record -- but only where the reference looks it up: on_match2 returns before manager.find under -v and -c
// (main.cpp:274-283), hash_search always calls it (main.cpp:315, side effects included), so exact look-ups with
// files stay on the fields path when -v or -c is given.
enum RawMode { RAW_NO, RAW_PLAIN, RAW_ROUTED };
RawMode raw_mode(const Job &job, Strategy st, const char *file2)
{
    const char *mode = getenv("FXTRACT_PATH");
    if (mode && !strcmp(mode, "fields")) return RAW_NO;
    const CliOptions &o = job.opt;
    // mates: the device applies the strategy's pair rule (run_raw_pairs); -l keeps its per-pair quirks (which file
    // name is printed, main.cpp:427-435) on the fields path
    if ((file2 || o.I) && o.l) return RAW_NO;
    if ((file2 || o.I) && st == S_MULTI && o.Q) return RAW_NO;     // read 2 is searched in its SEQUENCE under -Q (main.cpp:416)
    if (job.router.all_stdout()) return RAW_PLAIN;   // includes -l: the first batch with a match ends the input
    if (o.l) return RAW_NO;                          // -l with output files: find()'s side effects need the pattern
    if (st == S_MULTI && (o.v || o.c)) return RAW_PLAIN;
    if ((st == S_MULTI || st == S_EXACT) && !o.v && !o.c) return RAW_ROUTED;
    return RAW_NO;
}

int reader_threads_from_env()
{
    const char *e = getenv("FXTRACT_READERS");
    int n = e ? atoi(e) : 0;
    if (n <= 0) { n = (int)std::thread::hardware_concurrency(); if (n > 16) n = 16; }
    return n < 1 ? 1 : n;
}

// Returns: -1 = raw path not applicable (nothing consumed), otherwise the byte offset from which
// the fields path has to continue (== file size when the raw path handled everything).
int64_t run_raw(Job &job, Strategy st, const char *file1, int fd, DevicePool &pool, Stats &stats)
{
    struct stat sb;
    if (fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) return -1;
    unsigned char magic[2] = {0, 0};
    if (pread(fd, magic, 2, 0) < 1 || magic[0] != '@') return -1;      // FASTA, gzip (1f 8b), empty, ...: general reader
    const CliOptions &o = job.opt;
    const uint64_t file_size = (uint64_t)sb.st_size;
    const size_t pad = std::min<size_t>((size_t)1 << 20, pool.batch_bytes() / 4) & ~(size_t)63;
    const size_t window = pool.batch_bytes() - pad;                    // a batch = carried tail + one window <= batch_bytes
    const int nbuf = (int)pool.capacity() + 3;                         // in flight on the devices + being cut + two being filled
    const double t_setup0 = now_s();
    RawIngest ingest(fd, file_size, window, pad, nbuf, reader_threads_from_env());
    stats.t_setup += now_s() - t_setup0;
    std::deque<RawPending> pending;
    int64_t reroute_at = -1;
    // -l (main.cpp:611-615, :439-444, :318-323): the first match prints the file name and ends the input; nothing else
    // is ever printed or counted.  simple / exact search only get there without -v (:608, :316), multi always (:439).
    const bool l_effective = o.l && (st == S_MULTI || !o.v);
    bool l_stop = false;
    auto finish_oldest = [&]() {
        fxg_result r;
        const double t0 = now_s();
        pool.collect(&r);
        const double t1 = now_s();
        stats.t_gpu += t1 - t0;
        RawPending p = pending.front();
        pending.pop_front();
        if (reroute_at < 0 && !l_stop) {
            if (!r.canonical_ok) { reroute_at = (int64_t)p.file_off; stats.rerouted++; }
            else {
                if (o.l) {
                    if (l_effective && r.n_matched > 0) { printf("%s\n", file1); l_stop = true; }
                } else if (o.c) job.matched_records += (long)r.n_selected;
                else if (r.run_pattern_id) {                          // per-pattern output files (main.cpp:285-287, :315-324)
                    for (uint64_t k = 0; k < r.n_runs; ++k) {
                        FILE *out = job.router.find(r.run_pattern_id[k]);
                        if (out) fwrite(p.host + r.runs[k].off, 1, r.runs[k].len, out);
                    }
                } else write_runs(p.host, r.runs, r.n_runs);
                stats.records += r.n_records;
                stats.t_write += now_s() - t1;
            }
        }                                                              // results after a rejected batch are discarded
        ingest.release_window(p.window);
    };
    const uint8_t *carry = nullptr;      // unconsumed tail of the previous window's batch (a partial record)
    size_t carry_n = 0;
    uint64_t consumed = 0;               // file offset up to which batches have been cut
    for (uint64_t k = 0; k < ingest.windows() && reroute_at < 0 && !l_stop; ++k) {
        if (pool.inflight() == pool.capacity()) { finish_oldest(); if (reroute_at >= 0 || l_stop) break; }
        size_t wlen;
        const double tw = now_s();
        uint8_t *w = ingest.wait_window(k, &wlen);
        stats.t_ingest += now_s() - tw;
        if (carry_n > pad) {                            // a record longer than the pad: not the fast path's business
            while (!pending.empty()) finish_oldest();
            if (reroute_at < 0) reroute_at = (int64_t)consumed;
            break;
        }
        uint8_t *batch = w - carry_n;
        if (carry_n) memcpy(batch, carry, carry_n);
        const size_t have = carry_n + wlen;
        const bool last = k + 1 == ingest.windows();
        size_t n = have;
        if (!last) n = fxg_find_record_cut(batch, have);
        if (!last && n == 0) {                          // no record boundary in a whole window
            while (!pending.empty()) finish_oldest();
            if (reroute_at < 0) reroute_at = (int64_t)consumed;
            break;
        }
        carry = batch + n; carry_n = have - n;
        const double ts = now_s();
        pool.submit_raw_from(batch, n, last);
        stats.t_submit += now_s() - ts;
        pending.push_back(RawPending{batch, consumed, n, k});
        consumed += n;
        stats.raw_bytes += n; stats.raw_batches++;
        // the carried tail lives in window k's buffer: it is copied into window k + 1's pad before window k can be
        // released, which happens only when batch k has been collected -- at least one loop iteration later
    }
    while (!pending.empty()) finish_oldest();
    if (l_stop) return (int64_t)file_size;
    if (reroute_at >= 0) {
        const char *mode = getenv("FXTRACT_PATH");
        if (mode && !strcmp(mode, "raw")) {
            fprintf(stderr, "fxtract: FXTRACT_PATH=raw but the input is not canonical FASTQ at byte %lld\n", (long long)reroute_at);
            exit(1);
        }
        return reroute_at;
    }
    return (int64_t)file_size;
}

// ------------------------------------------------------------------ raw path, mates (two files / -I)
// Fxstream::read (fx.cpp:58-77) pairs record i of file 1 with record i of file 2 (or records 2i, 2i+1 of one
// interleaved file) and ends at the first failed read of either.  Both streams are read in windows like run_raw;
// the reader threads also count the '\n' of every slice, so the host can cut both streams at the same number of
// whole records (four lines each -- verified on the device, which raises FXG_V_PAIRING otherwise) without parsing.
struct MateStream {
    std::unique_ptr<RawIngest> in;
    size_t pad = 0;
    const uint8_t *carry = nullptr; size_t carry_n = 0; uint64_t carry_lines = 0;
    uint64_t consumed = 0;                 // file offset where the next batch starts
    uint8_t *batch = nullptr; size_t have = 0; uint64_t lines = 0;
    // carried tail + window k -> batch / have / lines; false: the carry does not fit in front of the window
    bool load(uint64_t k, double &t_wait)
    {
        size_t wlen;
        const double t0 = now_s();
        uint8_t *w = in->wait_window(k, &wlen);
        t_wait += now_s() - t0;
        if (carry_n > pad) return false;
        batch = w - carry_n;
        if (carry_n) memmove(batch, carry, carry_n);
        have = carry_n + wlen;
        lines = carry_lines + in->window_lines(k);
        return true;
    }
    // keep the first `recs` records for this batch, carry the rest; returns the batch's byte length
    size_t cut(uint64_t recs)
    {
        const size_t n = recs ? offset_behind_newline_from_end(batch, have, lines - 4 * recs) : 0;
        carry = batch + n; carry_n = have - n; carry_lines = lines - 4 * recs;
        return n;
    }
};

struct MatePending { const uint8_t *host1, *host2; size_t n1; uint64_t off1, off2, window; };

// Returns false when the raw path is not applicable (nothing consumed); else `next1` / `next2` are the offsets from
// which the fields path has to continue (== file sizes when everything was handled here).
bool run_raw_pairs(Job &job, Strategy st, int fd1, int fd2, DevicePool &pool, Stats &stats, uint64_t &next1, uint64_t &next2, bool &finished)
{
    const bool two = fd2 >= 0;
    struct stat sb1, sb2;
    if (fstat(fd1, &sb1) != 0 || !S_ISREG(sb1.st_mode)) return false;
    if (two && (fstat(fd2, &sb2) != 0 || !S_ISREG(sb2.st_mode))) return false;
    unsigned char magic = 0;
    if (pread(fd1, &magic, 1, 0) < 1 || magic != '@') return false;
    if (two && (pread(fd2, &magic, 1, 0) < 1 || magic != '@')) return false;
    const CliOptions &o = job.opt;
    const uint64_t size1 = (uint64_t)sb1.st_size, size2 = two ? (uint64_t)sb2.st_size : 0;
    const size_t pad = std::min<size_t>((size_t)1 << 20, pool.batch_bytes() / 8) & ~(size_t)63;
    const int nstreams = two ? 2 : 1;
    if (pool.batch_bytes() < (size_t)nstreams * pad + 4096) return false;
    const size_t usable = pool.batch_bytes() - (size_t)nstreams * pad - 128;
    const uint64_t nwin = std::max<uint64_t>(1, (size1 + size2 + usable - 1) / usable);
    const int nbuf = (int)pool.capacity() + 3;
    const int readers = std::max(1, reader_threads_from_env() / nstreams);
    MateStream ms[2];
    const double t_setup0 = now_s();
    for (int i = 0; i < nstreams; ++i) {
        const uint64_t size = i ? size2 : size1;
        const size_t window = (size_t)((((size + nwin - 1) / nwin) + 63) & ~(uint64_t)63);
        ms[i].pad = pad;
        ms[i].in.reset(new RawIngest(i ? fd2 : fd1, size, std::max<size_t>(window, 64), pad, nbuf, readers, true, nwin));
    }
    stats.t_setup += now_s() - t_setup0;
    std::deque<MatePending> pending;
    bool reroute = false;
    uint64_t re1 = 0, re2 = 0;
    std::vector<struct iovec> iov;
    auto finish_oldest = [&]() {
        fxg_result r;
        const double t0 = now_s();
        pool.collect(&r);
        const double t1 = now_s();
        stats.t_gpu += t1 - t0;
        MatePending p = pending.front();
        pending.pop_front();
        if (!reroute) {
            if (!r.canonical_ok) { reroute = true; re1 = p.off1; re2 = p.off2; stats.rerouted++; }
            else {
                if (o.c) job.matched_records += (long)r.n_selected;
                else if (!two && !r.run_pattern_id) write_runs(p.host1, r.runs, r.n_runs);      // interleaved: a pair is one byte range
                else {
                    // one run per record: interleaved -- mates are runs 2k, 2k+1; two files -- runs k and n/2 + k,
                    // the second with its offset relative to the concatenation of both parts
                    const uint64_t np = r.n_runs / 2;
                    auto mate = [&](uint64_t k, int which, const uint8_t *&ptr, size_t &len, int &pid) {
                        const uint64_t idx = two ? (which ? np + k : k) : 2 * k + (uint64_t)which;
                        const fxg_run &run = r.runs[idx];
                        ptr = (two && which) ? p.host2 + (run.off - p.n1) : p.host1 + run.off;
                        len = (size_t)run.len;
                        pid = r.run_pattern_id ? r.run_pattern_id[idx] : -1;
                    };
                    if (r.run_pattern_id) {                                 // per-pattern output files: read 1's pattern if it
                        for (uint64_t k = 0; k < np; ++k) {                 // has one, else read 2's (main.cpp:408-435, :315-337)
                            const uint8_t *a, *b; size_t la, lb; int pa, pb;
                            mate(k, 0, a, la, pa); mate(k, 1, b, lb, pb);
                            FILE *out = job.router.find(pa >= 0 ? pa : pb);
                            if (out) { fwrite(a, 1, la, out); fwrite(b, 1, lb, out); }
                        }
                    } else {
                        fflush(stdout);
                        constexpr size_t kIov = 512;
                        for (uint64_t k0 = 0; k0 < np; k0 += kIov / 2) {
                            iov.clear();
                            size_t want = 0;
                            for (uint64_t k = k0; k < np && k < k0 + kIov / 2; ++k)
                                for (int which = 0; which < 2; ++which) {
                                    const uint8_t *ptr; size_t len; int pid;
                                    mate(k, which, ptr, len, pid);
                                    iov.push_back({const_cast<uint8_t *>(ptr), len});
                                    want += len;
                                }
                            size_t first = 0;
                            while (want) {
                                ssize_t got = writev(STDOUT_FILENO, iov.data() + first, (int)(iov.size() - first));
                                if (got < 0) { if (errno == EINTR) continue; perror("fxtract: write"); exit(1); }
                                want -= (size_t)got;
                                while (got > 0 && first < iov.size()) {
                                    if ((size_t)got >= iov[first].iov_len) { got -= (ssize_t)iov[first].iov_len; ++first; }
                                    else { iov[first].iov_base = (uint8_t *)iov[first].iov_base + got; iov[first].iov_len -= (size_t)got; got = 0; }
                                }
                            }
                        }
                    }
                }
                stats.records += r.n_records;
                stats.t_write += now_s() - t1;
            }
        }
        for (int i = 0; i < nstreams; ++i) ms[i].in->release_window(p.window);
    };
    auto drain = [&]() { while (!pending.empty()) finish_oldest(); };
    auto reroute_here = [&]() {
        drain();
        if (!reroute) { reroute = true; re1 = ms[0].consumed; re2 = ms[1].consumed; }
    };
    for (uint64_t k = 0; k < nwin && !reroute; ++k) {
        if (pool.inflight() == pool.capacity()) { finish_oldest(); if (reroute) break; }
        bool fits = true;
        for (int i = 0; i < nstreams; ++i) fits = ms[i].load(k, stats.t_ingest) && fits;
        if (!fits) { reroute_here(); break; }                       // a carried tail longer than the pad
        const bool last = k + 1 == nwin;
        if (last) {
            // a last line without '\n' or a record cut short is the tokenizer's business (kseq.hpp:120-164)
            bool whole = true;
            for (int i = 0; i < nstreams; ++i)
                if (ms[i].have && (ms[i].batch[ms[i].have - 1] != '\n' || (ms[i].lines & 3))) whole = false;
            if (!whole) { reroute_here(); break; }
        }
        uint64_t recs = ms[0].lines / 4;
        if (two) recs = std::min<uint64_t>(recs, ms[1].lines / 4);
        else recs &= ~1ull;                                          // whole pairs; an odd last record has no mate (fx.cpp:64-69)
        if (recs == 0) {
            if (last) break;                                         // one stream has no record left: the pair stream ends
            reroute_here();                                          // no whole pair in a whole window
            break;
        }
        const size_t n1 = ms[0].cut(recs), n2 = two ? ms[1].cut(recs) : 0;
        if (n1 == 0 || (two && n2 == 0)) { reroute_here(); break; }
        const double ts = now_s();
        pool.submit_raw_pairs_from(ms[0].batch, n1, two ? ms[1].batch : nullptr, n2, last);
        stats.t_submit += now_s() - ts;
        pending.push_back(MatePending{ms[0].batch, two ? ms[1].batch : nullptr, n1, ms[0].consumed, ms[1].consumed, k});
        ms[0].consumed += n1; ms[1].consumed += n2;
        stats.raw_bytes += n1 + n2; stats.raw_batches++;
    }
    drain();
    finished = !reroute;
    if (reroute) {
        const char *mode = getenv("FXTRACT_PATH");
        if (mode && !strcmp(mode, "raw")) {
            fprintf(stderr, "fxtract: FXTRACT_PATH=raw but the mates are not canonical FASTQ at bytes %llu / %llu\n",
                    (unsigned long long)re1, (unsigned long long)re2);
            exit(1);
        }
        next1 = re1; next2 = re2;
    } else { next1 = size1; next2 = size2; }
    return true;
}

// ------------------------------------------------------------------ POSIX regex (outside the graft; CPU, main.cpp:462-514)
int run_regex(Job &job, PairSource &src, bool paired, regex_t *re)
{
    const CliOptions &o = job.opt;
    std::string tmp;
    auto field_of = [&](const PairBatch &b, const RecView &r) -> const char * {
        const char *p; uint32_t n;
        target_slice(o, S_EXACT, b, r, false, p, n);
        tmp.assign(p, n);
        return tmp.c_str();
    };
    auto regfail = [&](int code) {
        size_t len = regerror(code, re, nullptr, 0);
        std::string msg(len, '\0');
        regerror(code, re, &msg[0], len);
        fprintf(stderr, "%s\n", msg.c_str());
    };
    for (;;) {
        PairBatch b;
        src.read_batch(b, 1u << 16, 32u << 20);
        for (size_t i = 0; i < b.size(); ++i) {
            int ret = regexec(re, field_of(b, b.first[i]), 0, nullptr, 0);
            if (ret == REG_NOMATCH) {
                if (paired && !b.second[i].empty()) {
                    ret = regexec(re, field_of(b, b.second[i]), 0, nullptr, 0);
                    if (ret == REG_NOMATCH && o.v) on_match(job, stdout, b, i);
                    if (ret != REG_NOMATCH && ret != 0) { regfail(ret); return 1; }
                    else if (ret == 0 && on_match(job, stdout, b, i) == 1) { printf("%s\n", src.name2().c_str()); return 0; }
                } else if (o.v) on_match(job, stdout, b, i);
            } else if (ret != 0) { regfail(ret); return 1; }
            else if (!o.v && on_match(job, stdout, b, i) == 1) { printf("%s\n", src.name1().c_str()); return 0; }
        }
        if (b.ended) break;
    }
    return 0;
}

}  // namespace

int run_input(Job &job, const char *file1, const char *file2)
{
    const CliOptions &o = job.opt;
    const bool paired = (file2 != nullptr) || o.I;
    PairSource src;
    if (src.open(file1, file2, o.I) != 0) {
        fprintf(stderr, "Failed to open stream\n");
        return 1;
    }
    if (o.E || o.G) {                                            // dispatch priority: regex first (main.cpp:707-725)
        regex_t re;
        int flags = REG_NOSUB | (o.E ? REG_EXTENDED : REG_BASIC);
        int rc = regcomp(&re, job.pattern.c_str(), flags);
        if (rc) {
            size_t len = regerror(rc, &re, nullptr, 0);
            std::string msg(len, '\0');
            regerror(rc, &re, &msg[0], len);
            fprintf(stderr, "%s\n", msg.c_str());
            return 1;
        }
        run_regex(job, src, paired, &re);
        regfree(&re);
        return 0;
    }
    const Strategy st = o.X ? S_EXACT : (o.f ? S_MULTI : S_LITERAL);      // X beats f (main.cpp:749-757)
    std::vector<std::string> pats;
    if (st == S_LITERAL) {
        pats.push_back(job.pattern);
        // prepare_revcomp (main.cpp:557-565): XOR rule; an empty reverse complement is not searched (:602)
        if (o.r != (o.H || o.Q || o.C)) {
            std::string rc = reverse_complement(job.pattern);
            if (!rc.empty()) pats.push_back(rc);
        }
    } else {
        pats = job.router.patterns();
        if (pats.empty()) {
            if (st == S_MULTI) fprintf(stderr, "problem initalizing multi-search\n");   // ssearch_create(…, 0) == NULL
            // exact lookup in an empty map never hits: only -v prints; handled below with a pattern that cannot occur
            if (st == S_MULTI) return 0;
            pats.push_back(std::string("\n\n", 2));
        }
    }
    PoolConfig pc;
    pc.mode = st == S_LITERAL ? FXG_MODE_LITERAL : (st == S_MULTI ? FXG_MODE_MULTI : FXG_MODE_EXACT);
    pc.target = fxg_target(o);
    pc.invert = o.v; pc.count_only = o.c || o.l; pc.want_pattern_id = st != S_LITERAL;      // -l needs no runs either
    RawMode rm = raw_mode(job, st, file2);
    pc.route_runs = st != S_LITERAL && !o.v && !o.c && !job.router.all_stdout();
    pc.batch_bytes = batch_bytes_from_env();
    pc.devices = devices_from_env();
    if (!job.pool) job.pool.reset(new DevicePool(pc, pats));     // strategy, flags and patterns are the same for every input
    DevicePool &pool = *job.pool;

    Stats stats;
    const auto t0 = std::chrono::steady_clock::now();
    uint64_t input_bytes = 0;
    bool done = false;
    if (rm == RAW_ROUTED && !pool.can_route()) rm = RAW_NO;       // pattern set not served by a fused matcher
    if (rm != RAW_NO && paired) {
        const int fd1 = open(file1, O_RDONLY), fd2 = file2 ? open(file2, O_RDONLY) : -1;
        if (fd1 >= 0 && (!file2 || fd2 >= 0)) {
            uint64_t next1 = 0, next2 = 0;
            bool finished = false;
            if (run_raw_pairs(job, st, fd1, fd2, pool, stats, next1, next2, finished)) {
                struct stat sb;
                fstat(fd1, &sb);
                input_bytes = (uint64_t)sb.st_size;
                if (fd2 >= 0) { fstat(fd2, &sb); input_bytes += (uint64_t)sb.st_size; }
                if (!finished) {
                    PairSource rest;                             // both streams stand at a FASTQ record boundary
                    rest.open_fds_at(fd1, next1, file1, fd2, next2, file2, o.I);
                    run_fields(job, st, rest, true, pool, stats);
                }
                done = true;
            }
        }
        if (fd1 >= 0) close(fd1);
        if (fd2 >= 0) close(fd2);
    } else if (rm != RAW_NO) {
        int fd = open(file1, O_RDONLY);
        if (fd >= 0) {
            int64_t next = run_raw(job, st, file1, fd, pool, stats);
            if (next >= 0) {
                struct stat sb;
                fstat(fd, &sb);
                input_bytes = (uint64_t)sb.st_size;
                if ((uint64_t)next >= (uint64_t)sb.st_size) done = true;
                else {
                    PairSource rest;                             // kseq has last_char == 0 at a FASTQ record boundary
                    rest.open_fd_at(fd, (uint64_t)next, file1);
                    run_fields(job, st, rest, false, pool, stats);
                    done = true;
                }
            }
            close(fd);
        }
    }
    if (!done) run_fields(job, st, src, paired, pool, stats);
    if (job.stats) {
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (stats.raw_batches)
            fprintf(stderr, "[fxtract] raw path main thread: %.3f s waiting for the readers, %.3f s for the devices, %.3f s writing, "
                            "%.3f s in submit calls, %.3f s allocating pinned windows\n",
                    stats.t_ingest, stats.t_gpu, stats.t_write, stats.t_submit, stats.t_setup);
        fprintf(stderr, "[fxtract] %s: raw %llu B in %llu batches, fields %llu B in %llu batches, %llu records, %llu re-routed, "
                        "%.3f s%s\n", file1, (unsigned long long)stats.raw_bytes, (unsigned long long)stats.raw_batches,
                (unsigned long long)stats.field_bytes, (unsigned long long)stats.field_batches,
                (unsigned long long)stats.records, (unsigned long long)stats.rerouted, s,
                input_bytes ? (", " + std::to_string(input_bytes / s / 1e9) + " GB/s").c_str() : "");
    }
    return 0;
}

}  // namespace fxh
