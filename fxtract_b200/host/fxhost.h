// fxhost.h -- host side of the fxtract drop-in CLI (C++17).
//
// The host keeps what north_star says it keeps -- a reader with kseq semantics, the
// FileManager output logic, the getopt surface -- and drives every matcher on the GPU
// through the C ABI of include/fxgpu.h.  Nothing in here matches patterns itself
// (the POSIX-regex strategies, which are outside the graft, are the one exception and
// stay on the CPU like in the reference).
#pragma once
#include <cstdint>
#include <cstdio>
#include <deque>
#include <map>
#include <memory>
#include <string>
#include <vector>
#include <zlib.h>
#include "fxgpu.h"

namespace fxh {

// ------------------------------------------------------------------ options (main.cpp:36-76, :110-198)
struct CliOptions {
    bool H = false, Q = false, C = false, G = false, E = false, X = false, r = false, I = false, P = false;
    bool v = false, one = false, two = false, S = false, c = false, l = false, L = false;
    const char *f = nullptr;
    int first_operand = 0;          // optind after parsing
};
[[noreturn]] void print_usage_and_exit();
CliOptions parse_cli(int argc, char **argv);
// util.cpp:3-17 with the table of util.h:11-20
std::string reverse_complement(const std::string &s);

// ------------------------------------------------------------------ reader (kseq.hpp:101-245 semantics)
// The reference tokenizer pulls bytes through a 4096-byte window; a handful of its
// edge behaviours (EOF detection = short refill, getuntil's early -1) depend on where
// the window ends, so the window is modelled exactly -- as a view that slides over a
// large read-ahead buffer instead of a 4 KiB copy.
class WindowedInput {
public:
    ~WindowedInput();
    void open_path(const char *path);                 // gzopen(path,"r"); failure is NOT an error (fx.cpp:32-33)
    void open_fd_at(int fd, uint64_t offset);         // raw-path re-route: continue at a byte offset
    int getc();                                       // kseq.hpp:167-181 incl. the signed-char cast
    // kseq.hpp:183-245; delimiter 0 = isspace, '\n' = newline.  Appends to `out`.
    // Returns -1 when called at end of input, else the number of bytes appended.
    long getuntil(int delimiter, std::vector<char> &out, int *dret);
    // bulk helpers used by the record parser (same consumption order as getc loops)
    const uint8_t *win_ptr() const { return store_.data() + wb_; }
    size_t win_left() const { return we_ - wb_; }
    void win_consume(size_t n) { wb_ += n; }
    bool win_refill_if_empty();                       // false = a getc() here would return -1 for end of input
private:
    bool refill();
    gzFile gz_ = nullptr;
    std::vector<uint8_t> store_;
    size_t hi_ = 0;                 // valid bytes in store_: [0, hi_)
    size_t wb_ = 0, we_ = 0;        // current 4096-byte window [wb_, we_)
    bool src_eof_ = false;          // underlying stream exhausted
    bool is_eof_ = false;           // kstream::is_eof: the last refill was short
};

enum Field { F_NAME = 0, F_COMMENT = 1, F_SEQ = 2, F_QUAL = 3 };

struct RecView {                    // one parsed record: four slices of the batch arena
    uint32_t off[4] = {0, 0, 0, 0};
    uint32_t len[4] = {0, 0, 0, 0};
    bool empty() const { return !len[0] && !len[1] && !len[2] && !len[3]; }      // kseq.cpp:13-15
    bool is_fasta() const { return len[F_QUAL] == 0; }                           // kseq.cpp:22-24
};

struct PairBatch {
    std::vector<char> arena;
    std::vector<RecView> first, second;     // same length; second[i].empty() when there is no mate
    bool ended = false;                     // the pair stream is exhausted after this batch
    void clear() { arena.clear(); first.clear(); second.clear(); ended = false; }
    size_t size() const { return first.size(); }
};

// fx.h:8-21 + fx.cpp:30-81: one or two inputs, optional interleaving.  `last_char` lives in the
// record slot, not in the stream (kseq.hpp:95), which is observable with -I on FASTA.
class PairSource {
public:
    // returns non-zero like Fxstream::open (second file + interleaved)
    int open(const char *file1, const char *file2, bool interleaved);
    void open_fd_at(int fd, uint64_t offset, const char *name);
    void open_fds_at(int fd1, uint64_t off1, const char *name1, int fd2, uint64_t off2, const char *name2, bool interleaved);
    // appends pairs until `max_pairs` / `max_arena` is reached or the stream ends
    void read_batch(PairBatch &b, size_t max_pairs, size_t max_arena);
    const std::string &name1() const { return name1_; }
    const std::string &name2() const { return name2_; }
private:
    int read_record(WindowedInput &in, int &last_char, std::vector<char> &arena, RecView &r);
    WindowedInput in1_, in2_;
    bool have2_ = false, interleaved_ = false, done_ = false;
    int last_char_[2] = {0, 0};
    std::string name1_, name2_;
};

// kseq.cpp:32-48 (fprintf("%s") semantics: every field stops at its first NUL)
void print_record(FILE *out, const PairBatch &b, const RecView &r);
inline void print_pair(FILE *out, const PairBatch &b, size_t i)
{
    print_record(out, b, b.first[i]);
    print_record(out, b, b.second[i]);
}

// ------------------------------------------------------------------ output routing (fileManager.cpp)
class OutputRouter {
public:
    ~OutputRouter();
    void add(const std::string &pattern);                               // fileManager.cpp:123-169
    void add(const std::string &pattern, const std::string &filename);  // fileManager.cpp:76-120
    // sorted pattern list == std::map iteration order (main.cpp:358-360); ids index into it
    const std::vector<std::string> &patterns();
    FILE *find(int pattern_id);                                         // fileManager.cpp:171-202 (side effects kept)
    bool all_stdout() const;
    size_t size() const { return by_pattern_.size(); }
private:
    struct Sink { std::string path; FILE *fp = nullptr; bool open = false, is_stdout = false; size_t written = 0; };
    FILE *acquire(Sink &s);
    std::map<std::string, int> by_pattern_, by_file_;
    std::vector<Sink> sinks_;
    std::vector<std::string> sorted_;
    std::vector<int> sink_of_;
    bool sorted_valid_ = false;
    int open_count_ = 0;
};

// ------------------------------------------------------------------ device pool over the C ABI
struct PoolConfig {
    int mode = FXG_MODE_LITERAL, target = FXG_TARGET_SEQ;
    bool invert = false, count_only = false, want_pattern_id = false;
    bool route_runs = false;        // raw path: per-pattern output files (every selected record its own run + pattern id)
    size_t batch_bytes = 64u << 20;
    std::vector<int> devices;       // CUDA ordinals; batches go round-robin, results come back in order
};

class DevicePool {
public:
    DevicePool(const PoolConfig &cfg, const std::vector<std::string> &patterns);
    ~DevicePool();
    size_t capacity() const { return ctx_.size() * 2; }
    size_t inflight() const { return order_.size(); }
    size_t batch_bytes() const { return cfg_.batch_bytes; }
    bool can_route() const;                                     // the raw path can name the pattern of every selected record
    void acquire(void **pinned, size_t *cap);                   // next device in turn
    void submit_fields(size_t nbytes, const uint32_t *off, uint32_t n);
    void submit_raw(size_t nbytes, bool is_last);
    void submit_raw_from(const void *host, size_t nbytes, bool is_last);   // caller-owned pinned memory, next device in turn
    // mates on the raw path: host2 == nullptr: interleaved (records 2i, 2i+1); else n1 bytes of file 1 + n2 bytes of file 2
    void submit_raw_pairs_from(const void *host1, size_t n1, const void *host2, size_t n2, bool is_last);
    void collect(fxg_result *res);                              // oldest submitted batch
    void grow(size_t batch_bytes);                              // re-create the contexts with larger buffers (idle only)
private:
    void create();
    void destroy();
    [[noreturn]] void fail(fxg_ctx *c, const char *what);
    PoolConfig cfg_;
    std::vector<std::string> patterns_;
    std::vector<fxg_ctx *> ctx_;
    std::deque<int> order_;
    int turn_ = 0, lent_ = -1;
};

std::vector<int> devices_from_env();
size_t batch_bytes_from_env();

// ------------------------------------------------------------------ strategies (main.cpp:306-452, :596-652)
struct Job {
    CliOptions opt;
    OutputRouter router;
    std::string pattern;            // single command-line pattern (main.cpp:657)
    long matched_records = 0;       // main.cpp:80 (cumulative over inputs)
    bool stats = false;
    std::unique_ptr<DevicePool> pool;   // one set of device contexts (pattern tables, staging) for all inputs
};
// One input (pair of) file(s): picks the strategy (main.cpp:707-757) and runs it.  Returns the
// process exit status contribution (0 ok).
int run_input(Job &job, const char *file1, const char *file2);

}  // namespace fxh
