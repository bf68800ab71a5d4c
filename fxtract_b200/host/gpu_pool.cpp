// gpu_pool.cpp -- the host's only door to the GPU: libfxgpu's C ABI (include/fxgpu.h).
// One fxg_ctx per CUDA device; batches are dealt round-robin and collected in submission
// order, so emission stays in input order however many GPUs share the work
// (contiguous ranges, no data-path collective).
// Any device error is fatal and loud: there is no CPU matcher to fall back to.
#include "fxhost.h"
#include <cstdlib>

namespace fxh {

DevicePool::DevicePool(const PoolConfig &cfg, const std::vector<std::string> &patterns) : cfg_(cfg), patterns_(patterns)
{
    if (cfg_.devices.empty()) cfg_.devices.push_back(0);
    create();
}

DevicePool::~DevicePool() { destroy(); }

void DevicePool::fail(fxg_ctx *c, const char *what)
{
    fprintf(stderr, "fxtract: %s: %s\n", what, fxg_last_error(c));
    exit(1);
}

void DevicePool::create()
{
    std::string blob;
    std::vector<uint32_t> off(1, 0);
    for (const std::string &p : patterns_) { blob += p; off.push_back((uint32_t)blob.size()); }
    for (int dev : cfg_.devices) {
        fxg_config c = {};
        c.device = dev; c.mode = cfg_.mode; c.target = cfg_.target;
        c.invert = cfg_.invert; c.count_only = cfg_.count_only; c.want_pattern_id = cfg_.want_pattern_id;
        c.batch_bytes = cfg_.batch_bytes;
        c.route_runs = cfg_.route_runs;
        fxg_ctx *ctx = nullptr;
        if (fxg_create(&c, &ctx) != FXG_OK) fail(nullptr, "cannot open the GPU search path");
        if (fxg_set_patterns(ctx, (const uint8_t *)blob.data(), off.data(), (uint32_t)patterns_.size()) != FXG_OK)
            fail(ctx, "cannot load the pattern set");
        ctx_.push_back(ctx);
    }
    turn_ = 0; lent_ = -1;
}

bool DevicePool::can_route() const
{
    for (fxg_ctx *c : ctx_) if (!fxg_can_route(c)) return false;
    return !ctx_.empty();
}

void DevicePool::destroy()
{
    for (fxg_ctx *c : ctx_) fxg_destroy(c);
    ctx_.clear();
    order_.clear();
}

void DevicePool::grow(size_t batch_bytes)
{
    if (batch_bytes <= cfg_.batch_bytes) return;
    if (!order_.empty() || lent_ >= 0) { fprintf(stderr, "fxtract: internal error: grow() while busy\n"); exit(1); }
    destroy();
    cfg_.batch_bytes = batch_bytes;
    create();
}

void DevicePool::acquire(void **pinned, size_t *cap)
{
    fxg_ctx *c = ctx_[turn_];
    if (fxg_acquire_buffer(c, pinned, cap) != FXG_OK) fail(c, "fxg_acquire_buffer");
    lent_ = turn_;
}

void DevicePool::submit_fields(size_t nbytes, const uint32_t *off, uint32_t n)
{
    fxg_ctx *c = ctx_[lent_];
    if (fxg_submit_fields(c, nbytes, off, n) != FXG_OK) fail(c, "fxg_submit_fields");
    order_.push_back(lent_);
    turn_ = (lent_ + 1) % (int)ctx_.size();
    lent_ = -1;
}

void DevicePool::submit_raw(size_t nbytes, bool is_last)
{
    fxg_ctx *c = ctx_[lent_];
    if (fxg_submit_raw(c, nbytes, is_last ? 1 : 0) != FXG_OK) fail(c, "fxg_submit_raw");
    order_.push_back(lent_);
    turn_ = (lent_ + 1) % (int)ctx_.size();
    lent_ = -1;
}

void DevicePool::submit_raw_from(const void *host, size_t nbytes, bool is_last)
{
    fxg_ctx *c = ctx_[turn_];
    if (fxg_submit_raw_from(c, host, nbytes, is_last ? 1 : 0) != FXG_OK) fail(c, "fxg_submit_raw_from");
    order_.push_back(turn_);
    turn_ = (turn_ + 1) % (int)ctx_.size();
}

void DevicePool::submit_raw_pairs_from(const void *host1, size_t n1, const void *host2, size_t n2, bool is_last)
{
    fxg_ctx *c = ctx_[turn_];
    const int rc = host2 ? fxg_submit_raw_pair_from(c, host1, n1, host2, n2, is_last ? 1 : 0)
                         : fxg_submit_raw_interleaved_from(c, host1, n1, is_last ? 1 : 0);
    if (rc != FXG_OK) fail(c, "fxg_submit_raw_pair_from");
    order_.push_back(turn_);
    turn_ = (turn_ + 1) % (int)ctx_.size();
}

void DevicePool::collect(fxg_result *res)
{
    if (order_.empty()) { fprintf(stderr, "fxtract: internal error: collect() with nothing in flight\n"); exit(1); }
    fxg_ctx *c = ctx_[order_.front()];
    order_.pop_front();
    if (fxg_collect(c, res) != FXG_OK) fail(c, "fxg_collect");
}

}  // namespace fxh
