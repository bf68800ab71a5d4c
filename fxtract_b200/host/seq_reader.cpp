// seq_reader.cpp -- FASTA/FASTQ tokenizer with the reference's kseq semantics, batch oriented.
//
// Behaviour follows kstream::read / getc / getuntil (kseq.hpp:120-245) and Fxstream::read
// (fx.cpp:58-81); the implementation is new: records are parsed straight into a flat batch
// arena (four slices per record) so that target fields can be packed for the GPU and printed
// without per-record allocations, and the 4096-byte refill window of the reference is a
// sliding view over a 1 MiB read-ahead buffer.
#include "fxhost.h"
#include <cstring>
#include <unistd.h>

namespace fxh {

namespace {
constexpr size_t kWindow = 4096;            // kseq.hpp:109 buffer size (observable, see fxhost.h)
constexpr size_t kReadAhead = 1u << 20;

// byte classes of the sequence loop (kseq.hpp:138-144), indexed by the raw byte:
//   2 = loop terminator ('>' '+' '@' and 0xFF, which getc() returns as -1), 1 = isgraph -> kept, 0 = dropped
struct SeqClass {
    uint8_t t[256];
    SeqClass()
    {
        for (int b = 0; b < 256; ++b) t[b] = (b >= 33 && b <= 126) ? 1 : 0;
        t['>'] = t['+'] = t['@'] = 2;
        t[0xFF] = 2;
    }
};
const SeqClass kSeqClass;

inline bool c_isspace(int c) { return c == ' ' || (c >= 9 && c <= 13); }
}  // namespace

WindowedInput::~WindowedInput()
{
    if (gz_) gzclose(gz_);
}

void WindowedInput::open_path(const char *path)
{
    gz_ = gzopen(path, "r");
    if (gz_) gzbuffer(gz_, 1u << 18);
    store_.resize(kReadAhead + kWindow);
    hi_ = wb_ = we_ = 0;
    src_eof_ = (gz_ == nullptr);            // unopened input reads as empty (App. A #18)
    is_eof_ = false;
}

void WindowedInput::open_fd_at(int fd, uint64_t offset)
{
    int d = dup(fd);
    if (d >= 0 && lseek(d, (off_t)offset, SEEK_SET) == (off_t)offset) gz_ = gzdopen(d, "r");
    else if (d >= 0) close(d);
    if (gz_) gzbuffer(gz_, 1u << 18);
    store_.resize(kReadAhead + kWindow);
    hi_ = wb_ = we_ = 0;
    src_eof_ = (gz_ == nullptr);
    is_eof_ = false;
}

// Slides the window to the next <= 4096 bytes of the stream.  Sets is_eof_ exactly when the
// reference's refill would (fewer than 4096 bytes delivered).  Returns false if 0 bytes.
bool WindowedInput::refill()
{
    if (hi_ - we_ < kWindow && !src_eof_) {
        size_t keep = hi_ - we_;
        if (keep) memmove(store_.data(), store_.data() + we_, keep);
        hi_ = keep; we_ = 0;
        while (hi_ < kReadAhead && !src_eof_) {
            int want = (int)(kReadAhead + kWindow - hi_ > (1u << 30) ? (1u << 30) : kReadAhead + kWindow - hi_);
            int got = gzread(gz_, store_.data() + hi_, (unsigned)want);
            if (got <= 0) { src_eof_ = true; break; }
            hi_ += (size_t)got;
            if (got < want) src_eof_ = true;
        }
    }
    size_t n = hi_ - we_;
    if (n > kWindow) n = kWindow;
    wb_ = we_;
    we_ += n;
    if (n < kWindow) is_eof_ = true;
    return n != 0;
}

bool WindowedInput::win_refill_if_empty()
{
    if (wb_ < we_) return true;
    if (is_eof_) return false;
    return refill();
}

int WindowedInput::getc()
{
    if (!win_refill_if_empty()) return -1;
    return (int)(signed char)store_[wb_++];
}

long WindowedInput::getuntil(int delimiter, std::vector<char> &out, int *dret)
{
    if (dret) *dret = 0;
    if (wb_ >= we_ && is_eof_) return -1;
    long appended = 0;
    for (;;) {
        if (wb_ >= we_) {
            if (is_eof_) break;
            if (!refill()) break;
        }
        const uint8_t *p = store_.data() + wb_;
        const size_t n = we_ - wb_;
        size_t i = 0;
        if (delimiter > 1) {
            const void *hit = memchr(p, delimiter, n);
            i = hit ? (size_t)((const uint8_t *)hit - p) : n;
        } else {
            while (i < n && !c_isspace((int)(signed char)p[i])) ++i;
        }
        out.insert(out.end(), (const char *)p, (const char *)p + i);
        appended += (long)i;
        wb_ += i + 1;                       // may step one past the window, exactly like `begin = i + 1`
        if (i < n) {
            if (dret) *dret = (int)(signed char)p[i];
            break;
        }
        wb_ = we_;
    }
    return appended;
}

// kstream::read (kseq.hpp:120-164).  Returns >= 0 on success, -1 / -2 like the reference.
int PairSource::read_record(WindowedInput &in, int &last_char, std::vector<char> &arena, RecView &r)
{
    int c;
    if (last_char == 0) {
        for (;;) {                          // skip to the next '>' or '@'
            if (!in.win_refill_if_empty()) return -1;
            const uint8_t *p = in.win_ptr();
            const size_t n = in.win_left();
            size_t i = 0;
            while (i < n && p[i] != '>' && p[i] != '@' && p[i] != 0xFF) ++i;
            if (i < n) {
                c = (int)(signed char)p[i];
                in.win_consume(i + 1);
                break;
            }
            in.win_consume(n);
        }
        if (c == -1) return -1;
        last_char = c;
    }
    r = RecView();
    r.off[F_NAME] = (uint32_t)arena.size();
    long nl = in.getuntil(0, arena, &c);
    if (nl < 0) return -1;
    r.len[F_NAME] = (uint32_t)nl;
    r.off[F_COMMENT] = (uint32_t)arena.size();
    if (c != '\n') {
        long cl = in.getuntil('\n', arena, nullptr);
        r.len[F_COMMENT] = cl > 0 ? (uint32_t)cl : 0;
    }
    // sequence: every isgraph byte up to the next '>', '+' or '@' (anywhere, not only at line starts)
    r.off[F_SEQ] = (uint32_t)arena.size();
    c = -1;
    for (bool stop = false; !stop;) {
        if (!in.win_refill_if_empty()) { c = -1; break; }
        const uint8_t *p = in.win_ptr();
        const size_t n = in.win_left();
        size_t i = 0;
        while (i < n) {
            // run of kept bytes
            size_t j = i;
            while (j < n && kSeqClass.t[p[j]] == 1) ++j;
            if (j > i) arena.insert(arena.end(), (const char *)p + i, (const char *)p + j);
            i = j;
            if (i < n) {
                if (kSeqClass.t[p[i]] == 2) { c = (int)(signed char)p[i]; ++i; stop = true; break; }
                ++i;                        // dropped byte (newline, blank, control, high bit)
            }
        }
        in.win_consume(i);
    }
    r.len[F_SEQ] = (uint32_t)(arena.size() - r.off[F_SEQ]);
    if (c == '>' || c == '@') last_char = c;
    r.off[F_QUAL] = (uint32_t)arena.size();
    if (c != '+') return (int)r.len[F_SEQ];
    // rest of the '+' line
    for (;;) {
        c = in.getc();
        if (c == -1 || c == '\n') break;
        // fast-forward inside the window
        const uint8_t *p = in.win_ptr();
        const size_t n = in.win_left();
        size_t i = 0;
        while (i < n && p[i] != '\n' && p[i] != 0xFF) ++i;
        in.win_consume(i);
    }
    if (c == -1) return -2;
    // quality: next len(seq) bytes in 33..127, across lines; one more byte is consumed after it
    const uint32_t want = r.len[F_SEQ];
    uint32_t have = 0;
    while ((c = in.getc()) != -1 && have < want) {
        if (c >= 33 && c <= 127) { arena.push_back((char)c); ++have; }
        // bulk part
        const uint8_t *p = in.win_ptr();
        const size_t n = in.win_left();
        size_t i = 0;
        while (i < n && have < want) {
            const uint8_t b = p[i];
            if (b == 0xFF) break;
            if (b >= 33 && b <= 127) {
                size_t j = i + 1;
                const size_t lim = (size_t)(want - have) < n - i ? i + (want - have) : n;
                while (j < lim && p[j] >= 33 && p[j] <= 127) ++j;
                arena.insert(arena.end(), (const char *)p + i, (const char *)p + j);
                have += (uint32_t)(j - i);
                i = j;
            } else ++i;
        }
        in.win_consume(i);
    }
    r.len[F_QUAL] = have;
    last_char = 0;
    if (have != want) return -2;
    return (int)want;
}

int PairSource::open(const char *file1, const char *file2, bool interleaved)
{
    name1_ = file1 ? file1 : "";
    in1_.open_path(file1);
    if (file2 != nullptr && interleaved) {
        fprintf(stderr, "A second file cannot be given if the interleaved flag is set\n");
        return 1;
    } else if (file2 != nullptr) {
        name2_ = file2;
        in2_.open_path(file2);
        have2_ = true;
    }
    interleaved_ = interleaved;
    return 0;
}

void PairSource::open_fd_at(int fd, uint64_t offset, const char *name)
{
    name1_ = name ? name : "";
    in1_.open_fd_at(fd, offset);
}

// raw-path re-route of mates: continue at a record boundary of both streams (fd2 < 0: interleaved, one stream)
void PairSource::open_fds_at(int fd1, uint64_t off1, const char *name1, int fd2, uint64_t off2, const char *name2, bool interleaved)
{
    name1_ = name1 ? name1 : "";
    in1_.open_fd_at(fd1, off1);
    if (fd2 >= 0) {
        name2_ = name2 ? name2 : "";
        in2_.open_fd_at(fd2, off2);
        have2_ = true;
    }
    interleaved_ = interleaved;
}

void PairSource::read_batch(PairBatch &b, size_t max_pairs, size_t max_arena)
{
    while (!done_ && b.first.size() < max_pairs && b.arena.size() < max_arena) {
        const size_t mark = b.arena.size();
        RecView r1, r2;
        if (read_record(in1_, last_char_[0], b.arena, r1) < 0) { b.arena.resize(mark); done_ = true; break; }
        if (interleaved_) {
            if (read_record(in1_, last_char_[1], b.arena, r2) < 0) { b.arena.resize(mark); done_ = true; break; }
        } else if (have2_) {
            if (read_record(in2_, last_char_[1], b.arena, r2) < 0) { b.arena.resize(mark); done_ = true; break; }
        }
        b.first.push_back(r1);
        b.second.push_back(r2);
    }
    b.ended = done_;
}

static inline void put_field(FILE *out, const char *p, uint32_t n)
{
    const size_t m = strnlen(p, n);          // "%s": stop at the first NUL
    if (m) fwrite(p, 1, m, out);
}

void print_record(FILE *out, const PairBatch &b, const RecView &r)
{
    if (r.empty()) return;
    const char *a = b.arena.data();
    const bool fasta = r.is_fasta();
    fputc(fasta ? '>' : '@', out);
    put_field(out, a + r.off[F_NAME], r.len[F_NAME]);
    if (r.len[F_COMMENT]) {
        fputc(' ', out);
        put_field(out, a + r.off[F_COMMENT], r.len[F_COMMENT]);
    }
    fputc('\n', out);
    put_field(out, a + r.off[F_SEQ], r.len[F_SEQ]);
    if (fasta) { fputc('\n', out); return; }
    fwrite("\n+\n", 1, 3, out);
    put_field(out, a + r.off[F_QUAL], r.len[F_QUAL]);
    fputc('\n', out);
}

}  // namespace fxh
