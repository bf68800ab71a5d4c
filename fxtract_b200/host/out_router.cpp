// out_router.cpp -- pattern registry + output routing with FileManager's observable behaviour
// (fileManager.cpp:76-202): first mapping of a pattern wins; one sink per distinct file name
// ("" = stdout); files are opened lazily on the first lookup hit ("w", or "a" after a forced
// close) -- also under -c / -v, where nothing is written (App. A #10); at most 256 files stay
// open, the one with the fewest records goes first.
#include "fxhost.h"
#include <cstring>

namespace fxh {

namespace { constexpr int kMaxOpen = 256; }        // fileManager.h:24-26 OPEN_MAX

OutputRouter::~OutputRouter()
{
    for (Sink &s : sinks_)
        if (s.open && !s.is_stdout && s.fp) fclose(s.fp);
}

void OutputRouter::add(const std::string &pattern, const std::string &filename)
{
    if (by_pattern_.count(pattern)) return;
    auto it = by_file_.find(filename);
    int idx;
    if (it == by_file_.end()) {
        idx = (int)sinks_.size();
        Sink s;
        s.path = filename;
        sinks_.push_back(s);
        by_file_[filename] = idx;
    } else idx = it->second;
    by_pattern_[pattern] = idx;
    sorted_valid_ = false;
}

void OutputRouter::add(const std::string &pattern)
{
    if (by_pattern_.count(pattern)) return;
    auto it = by_file_.find("");
    int idx;
    if (it == by_file_.end()) {
        idx = (int)sinks_.size();
        Sink s;
        s.fp = stdout; s.open = true; s.is_stdout = true;
        sinks_.push_back(s);
        by_file_[""] = idx;
    } else idx = it->second;
    by_pattern_[pattern] = idx;
    sorted_valid_ = false;
}

const std::vector<std::string> &OutputRouter::patterns()
{
    if (!sorted_valid_) {
        sorted_.clear(); sink_of_.clear();
        for (const auto &kv : by_pattern_) { sorted_.push_back(kv.first); sink_of_.push_back(kv.second); }
        sorted_valid_ = true;
    }
    return sorted_;
}

bool OutputRouter::all_stdout() const
{
    for (const Sink &s : sinks_) if (!s.is_stdout) return false;
    return true;
}

FILE *OutputRouter::acquire(Sink &s)
{
    if (s.open) { ++s.written; return s.fp; }
    if (open_count_ >= kMaxOpen) {
        Sink *victim = nullptr;
        for (Sink &o : sinks_)
            if (o.open && !o.is_stdout && (!victim || o.written < victim->written)) victim = &o;
        if (victim) { fclose(victim->fp); victim->fp = nullptr; victim->open = false; --open_count_; }
    }
    s.fp = fopen(s.path.c_str(), s.written ? "a" : "w");
    if (!s.fp) { perror("problem opening file"); return nullptr; }
    s.open = true;
    if (!s.is_stdout) ++open_count_;
    ++s.written;
    return s.fp;
}

FILE *OutputRouter::find(int pattern_id)
{
    patterns();
    if (pattern_id < 0 || pattern_id >= (int)sink_of_.size()) return nullptr;
    return acquire(sinks_[sink_of_[pattern_id]]);
}

}  // namespace fxh
