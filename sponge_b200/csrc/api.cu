// C ABI of the B200 Poseidon sponge engine (include/poseidon_b200.h): contexts, the constant-bank
// slot allocator, host<->device staging pipeline, lock-step batches and the single-sponge drop-in.
// Host code only; every computation is a launch of the kernels registered by tu_sponge.cu /
// misc_kernels.cu.  There is deliberately no CPU implementation of the permutation in here.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/poseidon_b200.h"
#include "kernels.cuh"
#include "fp52_tables.hpp"
#include "tc_tables.hpp"
#include "opt_constants.hpp"
#include "params.hpp"
#include "registry.hpp"

namespace psd {

// ------------------------------------------------------------------------------------------------
// Errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string t_last_error;

static psd_status fail(psd_status s, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_last_error = buf;
    return s;
}

#define PSD_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e__ = (cudaError_t)(call);                                                      \
        if (e__ != cudaSuccess)                                                                     \
            return fail(PSD_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                                  \
    } while (0)

// ------------------------------------------------------------------------------------------------
// Registries filled by static initialisers of the kernel translation units
// ------------------------------------------------------------------------------------------------
struct PoolUploader {
    int field, t;
    PoolUploadFn fn;
};
static std::vector<PoolUploader>& pool_uploaders() {
    static std::vector<PoolUploader> v;
    return v;
}
static PoolUploadFn find_uploader(int field, int t) {
    for (const auto& u : pool_uploaders())
        if (u.field == field && u.t == t) return u.fn;
    return nullptr;
}
static std::vector<SpongeKernelEntry>& sponge_kernels() {
    static std::vector<SpongeKernelEntry> v;
    return v;
}
void register_pool_uploader(int field, int t, PoolUploadFn fn) { pool_uploaders().push_back(PoolUploader{field, t, fn}); }
void register_sponge_kernel(const SpongeKernelEntry& e) { sponge_kernels().push_back(e); }

// Best registered kernel for (field, t, alpha) of pool kind `pool` whose variant is <= `variant` and
// whose table requirements the descriptor meets (variant 3 reads the full digit tables, 4 the lane-update
// ones, 1-2 the regrouped constants, 5 the 52-bit-limb tables).
static const SpongeKernelEntry* find_kernel(int field, int t, uint64_t alpha, int variant, int pool,
                                            const PermDesc* d = nullptr, int ragged = 0) {
    const SpongeKernelEntry* best = nullptr;
    int best_score = -1;
    for (const auto& e : sponge_kernels()) {
        if (e.field != field || e.t != t || e.pool != pool || e.ragged != ragged) continue;
        if (e.alpha != 0 && e.alpha != alpha) continue;
        if (e.variant > variant) continue;
        if (d) {
            if (e.variant == 5 && !d->fp_ok) continue;
            if (e.variant == 3 && d->ctab_ok != 1) continue;
            if (e.variant == 4 && d->ctab_ok != 2) continue;
            if ((e.variant == 1 || e.variant == 2) && !d->opt_ok) continue;
        }
        const int score = (e.alpha == alpha ? 2 : 0) + e.variant * 4;
        if (score > best_score) {
            best_score = score;
            best = &e;
        }
    }
    return best;
}
static int kernel_index(const SpongeKernelEntry* k) { return (int)(k - sponge_kernels().data()); }

static std::atomic<uint64_t> g_launches{0};

// ------------------------------------------------------------------------------------------------
// Where a context's constants live on its device.  Two homes:
//   bank   a slot of the 62 KiB constant-bank pool of the kernel unit for (field, t) -- or of the
//          variant-5 unit, key kFp52PoolKey -- read by the kernels through uniform loads;
//   smem   a buffer in global memory that the shared-memory form of the kernels copies into each block's
//          shared memory; no size limit worth mentioning and no competition between parameter sets.
// Contexts whose images are byte-identical share one placement (reference pattern: many
// PoseidonSponge::new(&params) with the same params, mod.rs:219-230), so the bank only ever holds
// distinct parameter sets; when it is full the next set goes to the smem home instead of being refused.
// The small-kernel unit keeps [R, R^2] of both fields at fixed offsets, uploaded once per device.
// ------------------------------------------------------------------------------------------------
enum : int { kPoolBank = 0, kPoolSmem = 1, kPoolGlobal = 2 };   // kPoolGlobal: read in place (variant-6 tables)
// pool keys beyond the state widths 2..9
constexpr int kTcSmallKey = kFp52PoolKey + 1;   // variant 6: the few integer-side constants (staged to shared memory)
constexpr int kTcTabKey = kFp52PoolKey + 2;     // variant 6: the table image in global memory
constexpr int kTc5TabKey = kFp52PoolKey + 3;    // variant 7: the table image in the tcgen05 B-operand layout

struct Placement {
    int device = 0, field = 0, key = 0;
    int kind = kPoolBank;
    uint32_t off = 0;              // bank: word offset of the slot
    uint32_t* d_img = nullptr;     // smem: the image in global memory
    std::vector<uint32_t> words;   // the image (kept for the identity check)
    int refs = 0;
};
struct PoolSlot {
    uint32_t off, len;
};
struct DevicePool {
    std::mutex mu;
    std::vector<PoolSlot> used;    // sorted by off
    std::vector<Placement*> live;
};
static DevicePool g_pools[64][2][kTc5TabKey + 1];   // [..][kFp52PoolKey]: the FP64-pipe unit's pool; then variant 6's two
static std::once_flag g_misc_once[64];
static inline uint32_t misc_one_off(int field) { return (uint32_t)field * 16; }
static inline uint32_t misc_r2_off(int field) { return (uint32_t)field * 16 + 8; }

static bool slot_alloc(DevicePool& p, uint32_t len, uint32_t* off) {
    uint32_t cur = 0;
    size_t pos = 0;
    for (; pos < p.used.size(); pos++) {
        if (p.used[pos].off - cur >= len) break;
        cur = p.used[pos].off + p.used[pos].len;
    }
    if (cur + len > kConstPoolWords) return false;
    p.used.insert(p.used.begin() + pos, PoolSlot{cur, len});
    *off = cur;
    return true;
}
static void slot_free(DevicePool& p, uint32_t off) {
    for (size_t i = 0; i < p.used.size(); i++) {
        if (p.used[i].off == off) {
            p.used.erase(p.used.begin() + i);
            return;
        }
    }
}

// Finds or creates the placement of `words` (a multiple of 8 words) on `device`, which must be current.
// Tries, in order: an identical live image, a bank slot (if try_bank), a global-memory buffer (if
// try_smem).  The upload is complete when this returns.  nullptr: neither home was possible; *cerr holds
// the CUDA error if one occurred.
static Placement* placement_acquire(int device, int field, int key, const std::vector<uint32_t>& words, bool try_bank,
                                    bool try_smem, cudaError_t* cerr, bool global = false) {
    *cerr = cudaSuccess;
    DevicePool& p = g_pools[device][field][key];
    std::lock_guard<std::mutex> lk(p.mu);
    for (Placement* q : p.live) {
        if (q->words == words && ((q->kind == kPoolBank && try_bank) || (q->kind == kPoolSmem && try_smem) ||
                                  (q->kind == kPoolGlobal && global))) {
            q->refs++;
            return q;
        }
    }
    uint32_t off = 0;
    PoolUploadFn up = try_bank ? find_uploader(field, key) : nullptr;
    if (up && words.size() <= kConstPoolWords && slot_alloc(p, (uint32_t)words.size(), &off)) {
        cudaError_t e = (cudaError_t)up(off, words.data(), (uint32_t)words.size(), nullptr);
        if (e == cudaSuccess) e = cudaStreamSynchronize(nullptr);
        if (e != cudaSuccess) {
            slot_free(p, off);
            *cerr = e;
            return nullptr;
        }
        Placement* q = new Placement();
        q->device = device, q->field = field, q->key = key, q->kind = kPoolBank, q->off = off, q->words = words, q->refs = 1;
        p.live.push_back(q);
        return q;
    }
    if ((try_smem && words.size() <= kSmemPoolMaxWords) || global) {
        uint32_t* d = nullptr;
        cudaError_t e = cudaMalloc(&d, words.size() * 4);
        if (e == cudaSuccess) e = cudaMemcpy(d, words.data(), words.size() * 4, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            if (d) cudaFree(d);
            cudaGetLastError();
            *cerr = e;
            return nullptr;
        }
        Placement* q = new Placement();
        q->device = device, q->field = field, q->key = key, q->kind = global ? kPoolGlobal : kPoolSmem, q->d_img = d,
        q->words = words, q->refs = 1;
        p.live.push_back(q);
        return q;
    }
    return nullptr;
}

// Drops one reference; the last one frees the slot / buffer (device must be current for the smem home).
static void placement_release(Placement* q) {
    if (!q) return;
    DevicePool& p = g_pools[q->device][q->field][q->key];
    std::lock_guard<std::mutex> lk(p.mu);
    if (--q->refs > 0) return;
    for (size_t i = 0; i < p.live.size(); i++) {
        if (p.live[i] == q) {
            p.live.erase(p.live.begin() + i);
            break;
        }
    }
    if (q->kind == kPoolBank) slot_free(p, q->off);
    else if (q->d_img) cudaFree(q->d_img);
    delete q;
}

}  // namespace psd

using namespace psd;

// ------------------------------------------------------------------------------------------------
// Context
// ------------------------------------------------------------------------------------------------
static const int kStageSlots = 3;

struct psd_ctx {
    std::atomic<int> refs{1};   // the creator + every live psd_batch; torn down when it reaches 0
    int device = 0;
    int field = 0;
    const HostField* hf = nullptr;
    uint32_t full_rounds = 0, partial_rounds = 0, rate = 0, capacity = 0, t = 0;
    uint64_t alpha = 0;
    std::vector<Fe> ark_mont, mds_mont;   // Montgomery copies of the parameters
    PermDesc desc{};       // offsets into the home of the integer kernels' image (pl_int)
    PermDesc misc_desc{};  // same parameters with one/r2 offsets of the small-kernel unit's pool
    Placement* pl_int = nullptr;
    // variant 5 (tu_fp52.cu) has its own image and descriptor; fdesc.fp_ok == 0 when the context has none
    PermDesc fdesc{};
    Placement* pl_fp = nullptr;
    // variant 6 (tu_tc.cu): a small image of the integer-side constants (staged to shared memory) with its
    // descriptor, and the tensor-core table image in global memory; tcd.ok == 0 when the context has none
    PermDesc tdesc{};
    TcDesc tcd{};
    Placement* pl_tc = nullptr;
    Placement* pl_tctab = nullptr;
    // variant 7 (tu_tc5.cu): the same small image, its own table image (tcgen05 B-operand layout)
    TcDesc tcd5{};
    Placement* pl_tc5tab = nullptr;
    int variant = 7;   // degrades to 5 / 3 / 2 / 0 by what the context has constants and kernels for (resolve_kernel)
    int sm_count = 0;
    // staging for PSD_MEM_HOST calls
    std::mutex stage_mu;
    cudaStream_t streams[kStageSlots] = {nullptr, nullptr, nullptr};
    void* stage_in[kStageSlots] = {nullptr, nullptr, nullptr};
    void* stage_out[kStageSlots] = {nullptr, nullptr, nullptr};
    size_t stage_in_bytes = 0, stage_out_bytes = 0;
    int32_t* d_err = nullptr;   // non-canonical-input flag of the one-shot calls (batches have their own)
    int32_t* h_err = nullptr;   // pinned
    std::mutex err_mu;          // serialises readers of the flag pair above
    // psd_merkle_root work space, kept between calls (cudaMalloc/cudaFree of hundreds of MB per call
    // cost 5-25 ms and made the step time jitter); grown on demand, released with the context
    std::mutex merkle_mu;
    void* merkle_scratch = nullptr;
    size_t merkle_scratch_bytes = 0;
    void* merkle_leaves = nullptr;   // device copy of host leaves
    size_t merkle_leaves_bytes = 0;
    // psd_merkle_root_multi: on the first context of a call, the gathered subtree roots (this device's own
    // memory); on every context, the event that says "my root has arrived there"
    void* multi_roots = nullptr;
    cudaEvent_t multi_event = nullptr;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            prev = -1;
        }
        ok = cudaSetDevice(dev) == cudaSuccess;
        if (!ok) cudaGetLastError();
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

inline const uint32_t* as_u32(const uint64_t* p) { return reinterpret_cast<const uint32_t*>(p); }
inline uint32_t* as_u32(uint64_t* p) { return reinterpret_cast<uint32_t*>(p); }

bool valid_enum(int v, int max) { return v >= 0 && v <= max; }

// The host-side mirror of the in-kernel sequencer: only the bookkeeping, no field arithmetic
// (reference src/poseidon/mod.rs:121-182, 232-254, 320-342).
void advance_mode(uint32_t rate, uint32_t n_absorb, bool do_squeeze, uint32_t n_squeeze, int* mode,
                  uint32_t* index) {
    int m = *mode;
    uint32_t idx = *index;
    if (n_absorb) {
        if (m == kModeAbsorbing) {
            if (idx == rate) idx = 0;
        } else {
            idx = 0;
        }
        m = kModeAbsorbing;
        uint32_t rem = n_absorb;
        while (idx + rem > rate) {
            rem -= rate - idx;
            idx = 0;
        }
        idx += rem;
    }
    if (do_squeeze) {
        if (m == kModeAbsorbing || idx == rate) idx = 0;
        m = kModeSqueezing;
        uint32_t rem = n_squeeze;
        while (idx + rem > rate) {
            rem -= rate - idx;
            idx = 0;
        }
        idx += rem;
    }
    *mode = m;
    *index = idx;
}

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

LaunchCfg sponge_cfg(const psd_ctx* ctx, const SpongeKernelEntry* k, uint64_t n) {
    // One 128-thread block per 128 sponges, scheduled by the hardware: measured 11% faster on B200
    // than a persistent grid-stride grid (blocks start staggered, so warps sit in different phases of
    // the permutation and the multiplier pipe stays fed), and the kernel has no per-thread loop so
    // that ptxas keeps round constants on the uniform datapath.
    LaunchCfg cfg;
    cfg.block = 128;
    const uint64_t sms = (uint64_t)std::max(1, ctx->sm_count);
    if (k->pool == kPoolSmem) {
        // Shared-memory form: every block copies the constant image first, so blocks are as large as the
        // kernel's launch bounds allow once the batch fills the SMs, and shrink (in warps) for small batches
        // so that every SM still gets one.
        const uint64_t max_block = (uint64_t)k->max_block;
        const uint64_t per_sm = (n + sms - 1) / sms;
        cfg.block = (int)std::min<uint64_t>(max_block, std::max<uint64_t>(32, (per_sm + 31) / 32 * 32));
        if (k->variant == 7) {
            cfg.block = (cfg.block + 127) / 128 * 128;   // whole 128-sponge tiles
            static const int tiles_cap = env_int("PSD_TC5_TILES", 0);   // tuning runs
            if (tiles_cap > 0) cfg.block = std::min(cfg.block, tiles_cap * 128);
        }
    } else if (k->variant == 5) {
        // Variant 5's partial-round body (1.5 K instructions) does not fit the instruction caches when
        // the warps of a scheduler sit at different places in it.  Warps of ONE block start together and
        // stay close, so large blocks share their instruction fetches: measured on B200 (c2) 162.8 M
        // perm/s with 128-thread blocks, 172.9 M with 256, 175.0 M with 512 (one 128-register block per
        // SM).  Small batches keep smaller blocks so that every SM still gets one.
        cfg.block = n >= 512 * sms ? 512 : n >= 256 * sms ? 256 : 128;
        const int forced = env_int("PSD_FP52_BLOCK", 0);   // tuning runs
        if (forced >= 32 && forced <= 512 && forced % 32 == 0) cfg.block = forced;
    }
    cfg.grid = (int)std::max<uint64_t>(1, (n + cfg.block - 1) / cfg.block);
    return cfg;
}

LaunchCfg misc_cfg(const psd_ctx* ctx, uint64_t n) {
    LaunchCfg cfg;
    cfg.block = 256;
    const uint64_t full = (uint64_t)ctx->sm_count * 8;
    const uint64_t need = (n + cfg.block - 1) / cfg.block;
    cfg.grid = (int)std::max<uint64_t>(1, std::min(full, need));
    return cfg;
}

// What a launch on this context runs: the requested variant, degraded to what the context has constants
// and kernels for, together with the image (placement + descriptor) that kernel reads.
struct Resolved {
    const SpongeKernelEntry* k = nullptr;
    const Placement* pl = nullptr;
    const PermDesc* d = nullptr;
};
Resolved resolve_kernel(const psd_ctx* ctx, int ragged = 0) {
    Resolved r;
    // variant 6 uses warp-collective instructions, so it serves the lock-step calls only; ragged batches take
    // the next variant down
    if (ctx->variant >= 7 && ctx->pl_tc && ctx->pl_tc5tab && ctx->tcd5.ok) {
        // (ragged batches too, where tu_tc5.cu has the ragged form: its blocks permute while any sponge needs to)
        const SpongeKernelEntry* k = find_kernel(ctx->field, (int)ctx->t, ctx->alpha, 7, kPoolSmem, nullptr, ragged);
        if (k && k->variant == 7) {
            r.k = k, r.pl = ctx->pl_tc, r.d = &ctx->tdesc;
            return r;
        }
    }
    if ((ctx->variant == 6 || (ctx->variant > 6 && ctx->field == PSD_FIELD_BN254_FR)) && !ragged && ctx->pl_tc && ctx->pl_tctab &&
        ctx->tcd.ok) {
        const SpongeKernelEntry* k = find_kernel(ctx->field, (int)ctx->t, ctx->alpha, 6, kPoolSmem, nullptr, 0);
        if (k && k->variant == 6) {
            r.k = k, r.pl = ctx->pl_tc, r.d = &ctx->tdesc;
            return r;
        }
    }
    if (ctx->variant >= 5 && ctx->pl_fp && ctx->fdesc.fp_ok) {
        const SpongeKernelEntry* k = find_kernel(ctx->field, (int)ctx->t, ctx->alpha, 5, ctx->pl_fp->kind, &ctx->fdesc, ragged);
        if (k && k->variant == 5) {
            r.k = k, r.pl = ctx->pl_fp, r.d = &ctx->fdesc;
            return r;
        }
    }
    if (!ctx->pl_int) return r;
    const int want = std::min(ctx->variant, 4);
    r.k = find_kernel(ctx->field, (int)ctx->t, ctx->alpha, want, ctx->pl_int->kind, &ctx->desc, ragged);
    r.pl = ctx->pl_int, r.d = &ctx->desc;
    return r;
}

// Kernels of the shared-memory form need the opt-in shared-memory size; set once per (kernel, device).
std::atomic<uint64_t> g_smem_attr_done[512];

psd_status launch(psd_ctx* ctx, const SpongeArgs& a, cudaStream_t stream) {
    const Resolved r = resolve_kernel(ctx, a.in_offsets ? 1 : 0);
    const SpongeKernelEntry* k = r.k;
    if (!k) return fail(PSD_ERR_UNSUPPORTED, "no kernel for field %d, t=%u", ctx->field, ctx->t);
    if (a.n == 0) return PSD_OK;
    if (a.n > (uint64_t)0x7fffffff * 32) return fail(PSD_ERR_ARG, "batch too large for one launch (%llu sponges)", (unsigned long long)a.n);
    LaunchCfg cfg = sponge_cfg(ctx, k, a.n);
    SpongeArgs b = a;
    b.d = *r.d;   // same parameters, offsets into the home of this kernel's image
    size_t smem = 0;
    if (k->pool == kPoolSmem) {
        b.pool_img = r.pl->d_img;
        b.pool_words = (uint32_t)r.pl->words.size();
        smem = (size_t)b.pool_words * 4;
        if (k->variant == 6) {
            // + the per-warp scratch tiles behind the staged image
            b.tc = ctx->tcd;
            b.tc_tab = reinterpret_cast<const unsigned char*>(ctx->pl_tctab->d_img);
            static const int stagger = env_int("PSD_TC_STAGGER", 1500);   // cycles; measured on c2/c3: +0.5 % / +1.4 % over starting together
            b.tc_stagger = cfg.block >= 256 ? (uint32_t)std::max(0, stagger) : 0u;
            smem = ((smem + 127) & ~(size_t)127) + (size_t)(cfg.block / 32) * kTcWarpScratch;
        }
        if (k->variant == 7) {
            // + the block's static tables, the tiles' A slots and table double buffers, the tiles' mbarriers
            b.tc = ctx->tcd5;
            b.tc_tab = reinterpret_cast<const unsigned char*>(ctx->pl_tc5tab->d_img);
            smem = ((smem + 1023) & ~(size_t)1023) + tc5_smem_bytes(ctx->t, (uint32_t)cfg.block / 128);
        }
        const int ki = kernel_index(k);
        const uint64_t bit = 1ull << ctx->device;
        if (ki >= 0 && ki < 512 && !(g_smem_attr_done[ki].load(std::memory_order_acquire) & bit)) {
            PSD_CUDA(cudaFuncSetAttribute((const void*)k->fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(kSmemPoolMaxWords * 4)));
            g_smem_attr_done[ki].fetch_or(bit, std::memory_order_release);
        }
    }
    k->fn<<<cfg.grid, cfg.block, smem, stream>>>(b);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    PSD_CUDA(cudaGetLastError());
    return PSD_OK;
}

SpongeArgs base_args(const psd_ctx* ctx, uint64_t n) {
    SpongeArgs a;
    memset(&a, 0, sizeof a);
    a.d = ctx->desc;
    a.n = n;
    a.err = ctx->d_err;
    a.mode = kModeAbsorbing;
    return a;
}

// Reads and clears the non-canonical-input flag once `stream` is idle.
psd_status read_err_flag(int32_t* h_err, int32_t* d_err, cudaStream_t stream) {
    PSD_CUDA(cudaMemcpyAsync(h_err, d_err, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
    PSD_CUDA(cudaStreamSynchronize(stream));
    if (*h_err) {
        PSD_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int32_t), stream));
        PSD_CUDA(cudaStreamSynchronize(stream));
        return fail(PSD_ERR_NONCANONICAL, "a canonical input element was >= p");
    }
    return PSD_OK;
}
psd_status check_err_flag(psd_ctx* ctx, cudaStream_t stream) {
    std::lock_guard<std::mutex> lk(ctx->err_mu);
    return read_err_flag(ctx->h_err, ctx->d_err, stream);
}

psd_status ensure_staging(psd_ctx* ctx, size_t in_bytes, size_t out_bytes) {
    if (in_bytes > ctx->stage_in_bytes) {
        for (int s = 0; s < kStageSlots; s++) {
            if (ctx->stage_in[s]) cudaFree(ctx->stage_in[s]);
            ctx->stage_in[s] = nullptr;
        }
        ctx->stage_in_bytes = 0;
        for (int s = 0; s < kStageSlots; s++) PSD_CUDA(cudaMalloc(&ctx->stage_in[s], in_bytes));
        ctx->stage_in_bytes = in_bytes;
    }
    if (out_bytes > ctx->stage_out_bytes) {
        for (int s = 0; s < kStageSlots; s++) {
            if (ctx->stage_out[s]) cudaFree(ctx->stage_out[s]);
            ctx->stage_out[s] = nullptr;
        }
        ctx->stage_out_bytes = 0;
        for (int s = 0; s < kStageSlots; s++) PSD_CUDA(cudaMalloc(&ctx->stage_out[s], out_bytes));
        ctx->stage_out_bytes = out_bytes;
    }
    return PSD_OK;
}

// Copies rows [first, first+cnt) of a batch of n sponges with w elements each between host and a
// compact device buffer holding cnt sponges in the same layout.
psd_status copy_chunk(void* dst, const void* src, int layout, uint64_t n, uint64_t first, uint64_t cnt,
                      uint32_t w, bool to_device, cudaStream_t stream) {
    if (cnt == 0 || w == 0) return PSD_OK;
    const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    if (layout == kLayoutAos) {
        const size_t bytes = (size_t)cnt * w * 32;
        if (to_device)
            PSD_CUDA(cudaMemcpyAsync(dst, (const char*)src + (size_t)first * w * 32, bytes, kind, stream));
        else
            PSD_CUDA(cudaMemcpyAsync((char*)dst + (size_t)first * w * 32, src, bytes, kind, stream));
    } else {
        // 2*w rows of 16-byte words: host pitch n*16, device pitch cnt*16
        if (to_device)
            PSD_CUDA(cudaMemcpy2DAsync(dst, (size_t)cnt * 16, (const char*)src + (size_t)first * 16, (size_t)n * 16,
                                       (size_t)cnt * 16, (size_t)w * 2, kind, stream));
        else
            PSD_CUDA(cudaMemcpy2DAsync((char*)dst + (size_t)first * 16, (size_t)n * 16, src, (size_t)cnt * 16,
                                       (size_t)cnt * 16, (size_t)w * 2, kind, stream));
    }
    return PSD_OK;
}

// Shape of one staged (PSD_MEM_HOST) batch call.
struct StagedCall {
    uint64_t n;
    const void* h_in;
    uint32_t in_w;          // elements per sponge copied in
    void* h_out;
    uint32_t out_w;         // elements per sponge copied out (0 when out_bytes_per != 0)
    uint32_t out_bytes_per; // bytes per sponge copied out (squeeze_bytes)
    int layout;
    bool in_place;          // output overwrites the input staging buffer (permute_batch)
    // Only sponges [first0, first0 + count) of the n in the host buffers are processed (a shard of a
    // multi-device call); count == 0 means all of them.
    uint64_t first0 = 0, count = 0;
};

// H2D -> kernel -> D2H over kStageSlots streams; `fill` completes the SpongeArgs for a chunk.
template <class Fill>
psd_status run_staged(psd_ctx* ctx, const StagedCall& c, Fill fill) {
    if (c.n == 0) return PSD_OK;
    const uint64_t begin = c.first0, end = c.count ? c.first0 + c.count : c.n;
    if (begin >= end) return PSD_OK;
    std::lock_guard<std::mutex> lk(ctx->stage_mu);
    const size_t in_per = (size_t)c.in_w * 32;
    const size_t out_per = c.out_bytes_per ? c.out_bytes_per : (size_t)c.out_w * 32;
    // chunk: 2^16 sponges.  Three chunks in flight (one per stream) keep all SMs busy, and short chunks
    // shrink the un-overlapped first upload / last download: measured on C2 (2^24 states from pinned
    // memory) 2^16 -> 99.8 % of the device-resident rate, 2^20 -> 97.3 %.  With variant 5, whose large
    // blocks want at least 512 sponges per SM in a launch, 2^17 is the better chunk (172.4 M perm/s end
    // to end at 2^16, 174.4 M at 2^17, 173.6 M at 2^18, against 175.1 M device-resident).
    const size_t per = std::max<size_t>(1, in_per + (c.in_place ? 0 : out_per));
    static const int chunk_log2 = env_int("PSD_STAGE_CHUNK_LOG2", 17);   // development knob
    // Bounded by bytes first (256 MiB per slot); the 2^16-sponge floor that keeps launches large applies only
    // while a slot stays under 512 MiB, so wide inputs (c5: 32 KiB per sponge) get smaller chunks instead of
    // multi-GiB staging buffers.
    uint64_t chunk = std::max<uint64_t>((256ull << 20) / per, std::min<uint64_t>(1ull << 16, (512ull << 20) / per));
    chunk = std::max<uint64_t>(chunk, 1);
    chunk = std::min<uint64_t>(chunk, 1ull << chunk_log2);
    // Whole waves: a chunk that is a multiple of (SMs x sponges per block) leaves no partly filled last wave behind every
    // launch (2^17 sponges are 1.38 waves of variant 7's 640-sponge blocks).
    static const int whole_waves = env_int("PSD_STAGE_WAVES", 1);
    if (whole_waves) {
        const Resolved r = resolve_kernel(ctx);
        if (r.k) {
            const uint64_t wave = (uint64_t)sponge_cfg(ctx, r.k, chunk).block * (uint64_t)std::max(1, ctx->sm_count);
            if (chunk > wave) chunk = std::max<uint64_t>(1, (chunk + wave / 2) / wave) * wave;
        }
    }
    chunk = std::min<uint64_t>(chunk, end - begin);
    psd_status st = ensure_staging(ctx, std::max<size_t>(chunk * in_per, 16), c.in_place ? 0 : std::max<size_t>(chunk * out_per, 16));
    if (st != PSD_OK) return st;
    int slot = 0;
    for (uint64_t first = begin; first < end; first += chunk, slot = (slot + 1) % kStageSlots) {
        const uint64_t cnt = std::min<uint64_t>(chunk, end - first);
        cudaStream_t s = ctx->streams[slot];
        void* d_in = ctx->stage_in[slot];
        void* d_out = c.in_place ? d_in : ctx->stage_out[slot];
        st = copy_chunk(d_in, c.h_in, c.layout, c.n, first, cnt, c.in_w, true, s);
        if (st != PSD_OK) return st;
        SpongeArgs a = base_args(ctx, cnt);
        fill(a, d_in, d_out);
        st = launch(ctx, a, s);
        if (st != PSD_OK) return st;
        if (c.out_bytes_per) {
            PSD_CUDA(cudaMemcpyAsync((char*)c.h_out + (size_t)first * c.out_bytes_per, d_out,
                                     (size_t)cnt * c.out_bytes_per, cudaMemcpyDeviceToHost, s));
        } else {
            st = copy_chunk(c.h_out, d_out, c.layout, c.n, first, cnt, c.out_w, false, s);
            if (st != PSD_OK) return st;
        }
    }
    for (int s = 0; s < kStageSlots; s++) PSD_CUDA(cudaStreamSynchronize(ctx->streams[s]));
    return check_err_flag(ctx, ctx->streams[0]);
}

psd_status check_batch_args(const psd_ctx* ctx, const void* p, int layout, int repr, int mem) {
    if (!ctx) return fail(PSD_ERR_ARG, "null context");
    if (!p) return fail(PSD_ERR_ARG, "null buffer");
    if (!valid_enum(layout, 1) || !valid_enum(repr, 1) || !valid_enum(mem, 1))
        return fail(PSD_ERR_ARG, "bad layout/repr/mem value");
    return PSD_OK;
}

bool grow_device_buffer(void** buf, size_t* have, size_t need) {
    if (need <= *have) return true;
    if (*buf) cudaFree(*buf);
    *buf = nullptr;
    *have = 0;
    if (cudaMalloc(buf, need) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    *have = need;
    return true;
}

// Host leaves -> the context's device copy (kept between calls).  Caller holds merkle_mu.
psd_status merkle_upload_leaves(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, cudaStream_t s, const uint32_t** d_out) {
    if (!grow_device_buffer(&ctx->merkle_leaves, &ctx->merkle_leaves_bytes, n_leaves * 32))
        return fail(PSD_ERR_NOMEM, "cudaMalloc(%zu) for leaves failed", n_leaves * 32);
    if (cudaMemcpyAsync(ctx->merkle_leaves, leaves, n_leaves * 32, cudaMemcpyHostToDevice, s) != cudaSuccess)
        return fail(PSD_ERR_CUDA, "leaf upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    *d_out = static_cast<const uint32_t*>(ctx->merkle_leaves);
    return PSD_OK;
}

// Queues, on stream s, the reduction of n_leaves (a power of two >= 2) device leaves in src_repr to ONE
// element in dst_repr: every level is one launch of the sponge kernel over the pairs of the level below
// (absorb 2, squeeze 1), intermediate levels in Montgomery form in the context's scratch buffer
// ([level 1: n/2][level 2: n/4] elements, ping-pong above that; grown here, kept between calls).  The
// root goes to `dst` if given (device pointer), else into the scratch buffer; *where tells which.
// Caller holds ctx->merkle_mu and the device is current.  Nothing is synchronised here.
psd_status merkle_reduce_async(psd_ctx* ctx, const uint32_t* src, int src_repr, size_t n_leaves, uint32_t* dst,
                               int dst_repr, cudaStream_t s, const uint32_t** where) {
    const size_t half = n_leaves / 2, quarter = std::max<size_t>(n_leaves / 4, 1);
    if (!grow_device_buffer(&ctx->merkle_scratch, &ctx->merkle_scratch_bytes, (half + quarter) * 32))
        return fail(PSD_ERR_NOMEM, "cudaMalloc(%zu) for Merkle levels failed", (half + quarter) * 32);
    uint32_t* scratch = static_cast<uint32_t*>(ctx->merkle_scratch);
    uint32_t* bufs[2] = {scratch, scratch + half * 8};
    int which = 0;
    for (size_t pairs = half; pairs >= 1; pairs /= 2) {
        SpongeArgs a = base_args(ctx, pairs);
        a.in = src;
        a.n_absorb = 2;
        a.in_layout = kLayoutAos;
        a.in_repr = src_repr;
        a.do_squeeze = 1;
        a.n_squeeze = 1;
        a.out = (pairs == 1 && dst) ? dst : bufs[which];
        a.out_layout = kLayoutAos;
        a.out_repr = pairs == 1 ? dst_repr : kReprMontgomery;
        psd_status st = launch(ctx, a, s);
        if (st != PSD_OK) return st;
        src = a.out;
        src_repr = kReprMontgomery;
        which ^= 1;
        if (pairs == 1) break;
    }
    if (where) *where = src;
    return PSD_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// Library
// ------------------------------------------------------------------------------------------------
extern "C" {

const char* psd_version(void) { return "poseidon_b200 0.1.0 (sm_100a)"; }
const char* psd_last_error_string(void) { return t_last_error.c_str(); }
const char* psd_status_string(psd_status s) {
    switch (s) {
        case PSD_OK: return "PSD_OK";
        case PSD_ERR_ARG: return "PSD_ERR_ARG";
        case PSD_ERR_SHAPE: return "PSD_ERR_SHAPE";
        case PSD_ERR_NONCANONICAL: return "PSD_ERR_NONCANONICAL";
        case PSD_ERR_CUDA: return "PSD_ERR_CUDA";
        case PSD_ERR_UNSUPPORTED: return "PSD_ERR_UNSUPPORTED";
        case PSD_ERR_NOMEM: return "PSD_ERR_NOMEM";
    }
    return "PSD_ERR_?";
}
int psd_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
uint64_t psd_kernel_launch_count(void) { return g_launches.load(); }

// ------------------------------------------------------------------------------------------------
// Parameters
// ------------------------------------------------------------------------------------------------
psd_status psd_find_poseidon_ark_and_mds(int32_t field, uint64_t prime_bits, uint32_t rate, uint32_t full_rounds,
                                         uint32_t partial_rounds, uint32_t skip_matrices, uint64_t* ark,
                                         uint64_t* mds) {
    const HostField* hf = host_field(field);
    if (!hf) return fail(PSD_ERR_ARG, "unknown field %d", field);
    if (!ark || !mds) return fail(PSD_ERR_ARG, "null output");
    if (rate < 1 || rate > 64 || full_rounds + partial_rounds > 4096) return fail(PSD_ERR_ARG, "bad sizes");
    int rc = find_poseidon_ark_and_mds(*hf, prime_bits, rate, full_rounds, partial_rounds, skip_matrices,
                                       reinterpret_cast<Fe*>(ark), reinterpret_cast<Fe*>(mds));
    if (rc == -1) return fail(PSD_ERR_SHAPE, "prime_bits %llu != modulus size %d", (unsigned long long)prime_bits, hf->bits);
    if (rc != 0) return fail(PSD_ERR_ARG, "x_i + y_j = 0: not a Cauchy matrix");
    return PSD_OK;
}

psd_status psd_default_params(int32_t field, uint32_t rate, int optimized_for_weights, psd_params* out,
                              uint64_t* ark_buf, uint64_t* mds_buf) {
    const HostField* hf = host_field(field);
    if (!hf) return fail(PSD_ERR_ARG, "unknown field %d", field);
    if (!out || !ark_buf || !mds_buf) return fail(PSD_ERR_ARG, "null output");
    DefaultEntry e;
    if (!default_entry(field, rate, optimized_for_weights != 0, &e))
        return fail(PSD_ERR_UNSUPPORTED, "no default parameters for field %d rate %u", field, rate);
    psd_status st = psd_find_poseidon_ark_and_mds(field, (uint64_t)hf->bits, rate, e.full_rounds, e.partial_rounds,
                                                  e.skip_matrices, ark_buf, mds_buf);
    if (st != PSD_OK) return st;
    out->field = field;
    out->full_rounds = e.full_rounds;
    out->partial_rounds = e.partial_rounds;
    out->alpha = e.alpha;
    out->rate = rate;
    out->capacity = 1;   // traits.rs:96
    out->ark = ark_buf;
    out->mds = mds_buf;
    out->repr = PSD_REPR_CANONICAL;
    return PSD_OK;
}

psd_status psd_ctx_create(const psd_params* p, int device, psd_ctx** out) {
    if (!p || !out) return fail(PSD_ERR_ARG, "null argument");
    *out = nullptr;
    const HostField* hf = host_field(p->field);
    if (!hf) return fail(PSD_ERR_ARG, "unknown field %d", p->field);
    if (!valid_enum(p->repr, 1)) return fail(PSD_ERR_ARG, "bad repr");
    if (!p->ark || !p->mds) return fail(PSD_ERR_SHAPE, "ark/mds missing");
    const uint64_t t64 = (uint64_t)p->rate + p->capacity;
    const uint64_t rounds64 = (uint64_t)p->full_rounds + p->partial_rounds;
    if (p->rate == 0) return fail(PSD_ERR_SHAPE, "rate must be >= 1");
    if (t64 < 2 || t64 > 9) return fail(PSD_ERR_UNSUPPORTED, "state width %llu outside [2,9]", (unsigned long long)t64);
    if (rounds64 == 0 || rounds64 > 1024) return fail(PSD_ERR_UNSUPPORTED, "round count %llu outside [1,1024]", (unsigned long long)rounds64);
    const uint32_t t = (uint32_t)t64, rounds = (uint32_t)rounds64;
    int ndev = psd_device_count();
    if (device < 0 || device >= ndev || device >= 64)
        return fail(PSD_ERR_CUDA, "CUDA device %d not available (%d visible); this library has no CPU path", device, ndev);

    psd_ctx* ctx = new (std::nothrow) psd_ctx();
    if (!ctx) return fail(PSD_ERR_NOMEM, "out of host memory");
    ctx->device = device;
    ctx->field = p->field;
    ctx->hf = hf;
    ctx->full_rounds = p->full_rounds;
    ctx->partial_rounds = p->partial_rounds;
    ctx->rate = p->rate;
    ctx->capacity = p->capacity;
    ctx->t = t;
    ctx->alpha = p->alpha;
    // Default request: variant 7 (tcgen05 linear layer AND reductions; c2: 378 M perm/s against 220 M for variant 6,
    // c5: 380 M against 169 M for variant 3).  Where variant 7 has no kernel the request degrades: BN254 to 6, then 5;
    // BLS12-381 skips 6 (mma.sync form, S-box with full reductions: 126 M perm/s on c5, slower than the all-integer
    // variant 3) unless psd_ctx_set_variant(ctx, 6) asks for exactly that.
    ctx->variant = 7;

    // constants -> Montgomery
    const Fe* ark = reinterpret_cast<const Fe*>(p->ark);
    const Fe* mds = reinterpret_cast<const Fe*>(p->mds);
    ctx->ark_mont.resize((size_t)rounds * t);
    ctx->mds_mont.resize((size_t)t * t);
    for (size_t i = 0; i < ctx->ark_mont.size() + ctx->mds_mont.size(); i++) {
        const bool is_ark = i < ctx->ark_mont.size();
        const Fe& v = is_ark ? ark[i] : mds[i - ctx->ark_mont.size()];
        if (hf_cmp(v.v, hf->p) >= 0) {
            delete ctx;
            return fail(PSD_ERR_NONCANONICAL, "%s[%zu] >= p", is_ark ? "ark" : "mds", is_ark ? i : i - (size_t)rounds * t);
        }
        Fe m = p->repr == PSD_REPR_CANONICAL ? hf_to_mont(*hf, v) : v;
        if (is_ark) ctx->ark_mont[i] = m;
        else ctx->mds_mont[i - ctx->ark_mont.size()] = m;
    }

    // constant-bank image: [ark][mds][one][r2][optimised constants]
    std::vector<uint32_t> img;
    auto push_fe = [&img](const Fe& f) {
        for (int k = 0; k < 4; k++) {
            img.push_back((uint32_t)f.v[k]);
            img.push_back((uint32_t)(f.v[k] >> 32));
        }
    };
    PermDesc& d = ctx->desc;
    d.ark_off = (uint32_t)img.size();
    for (const Fe& f : ctx->ark_mont) push_fe(f);
    d.mds_off = (uint32_t)img.size();
    for (const Fe& f : ctx->mds_mont) push_fe(f);
    d.one_off = (uint32_t)img.size();
    push_fe(hf_one(*hf));
    d.r2_off = (uint32_t)img.size();
    Fe r2;
    memcpy(r2.v, hf->r2, 32);
    push_fe(r2);
    d.half_full = p->full_rounds / 2;
    d.partial = p->partial_rounds;
    d.total = rounds;
    d.rate = p->rate;
    d.capacity = p->capacity;
    d.alpha_lo = (uint32_t)p->alpha;
    d.alpha_hi = (uint32_t)(p->alpha >> 32);
    // optimised-form constants only where an optimised kernel exists (t <= 5)
    if (t <= 5)
        append_optimised_constants(*hf, t, p->full_rounds, p->partial_rounds, ctx->ark_mont, ctx->mds_mont, ctx->desc, img);

    // Digit tables for constant products (variant 3; t = 5: lane-update tables only, variant 4) are about 8x
    // the constants they cover.  The image WITH tables is tried first: an identical live image, a bank slot,
    // or -- when the bank is taken by other parameter sets -- the shared-memory home, whose kernels have
    // the same tables (PSD_PREFER_BANK=1 reverses the last two for measurements: bank without tables first).
    const size_t base_words = img.size();
    const PermDesc base_desc = ctx->desc;
    if (ctx->desc.opt_ok) {
        if (t == 3)
            append_digit_tables(*hf, t, p->full_rounds, p->partial_rounds, ctx->mds_mont, ctx->desc, img);
        else if (t == 5 && p->alpha == 5)   // the full set does not fit the bank at t = 5: lane-update tables only
            append_digit_tables(*hf, t, p->full_rounds, p->partial_rounds, ctx->mds_mont, ctx->desc, img, true);
    }
    while (img.size() % 8) img.push_back(0);

    DeviceGuard guard(device);
    if (!guard.ok) {
        delete ctx;
        return fail(PSD_ERR_CUDA, "cudaSetDevice(%d) failed", device);
    }
    {
        static const bool prefer_bank = env_int("PSD_PREFER_BANK", 0) != 0;
        static const bool force_smem = env_int("PSD_FORCE_SMEM", 0) != 0;   // test / measurement knob
        const bool has_tables = ctx->desc.ctab_ok != 0;
        cudaError_t cerr = cudaSuccess;
        if (has_tables) {
            const SpongeKernelEntry* sk = find_kernel(p->field, (int)t, p->alpha, 4, kPoolSmem, &ctx->desc);
            const bool smem_tables = sk && sk->variant >= 3;
            ctx->pl_int = placement_acquire(device, p->field, (int)t, img, !force_smem, smem_tables && !prefer_bank, &cerr);
        }
        if (!ctx->pl_int && cerr == cudaSuccess) {
            std::vector<uint32_t> base(img.begin(), img.begin() + base_words);
            while (base.size() % 8) base.push_back(0);
            ctx->pl_int = placement_acquire(device, p->field, (int)t, base, !force_smem, true, &cerr);
            if (ctx->pl_int) {
                ctx->desc = base_desc;
                img.swap(base);
            }
        }
        if (!ctx->pl_int) {
            delete ctx;
            if (cerr != cudaSuccess) return fail(PSD_ERR_CUDA, "uploading the constants failed: %s", cudaGetErrorString(cerr));
            return fail(PSD_ERR_NOMEM, "no home for the constant image (%zu words)", img.size());
        }
        if (ctx->pl_int->kind == kPoolBank) offset_desc(ctx->desc, ctx->pl_int->off);
    }

    // Variant 5 (FP64-pipe linear layer): its own image [ark of round 0][mds][one][r2][52-bit-limb tables],
    // for the parameter sets some unit has variant-5 kernels for.  Home: the pool of tu_fp52.cu when the
    // image fits and the pool has room (one t = 3 parameter set nearly fills it), else shared memory.
    {
        const SpongeKernelEntry* fk_bank = find_kernel(p->field, (int)t, p->alpha, 5, kPoolBank);
        const SpongeKernelEntry* fk_smem = find_kernel(p->field, (int)t, p->alpha, 5, kPoolSmem);
        const bool bank_ok = fk_bank && fk_bank->variant == 5 && fk_bank->alpha == p->alpha;
        const bool smem_ok = fk_smem && fk_smem->variant == 5 && fk_smem->alpha == p->alpha;
        if (bank_ok || smem_ok) {
            // base offsets are image-relative in base_desc
            std::vector<uint32_t> fimg;
            PermDesc fd = base_desc;
            fd.opt_ok = fd.ctab_ok = 0;
            fd.pre_mds_off = fd.sparse_off = fd.post_ark_off = 0;
            fd.ct_m00_off = fd.ct_mds_off = fd.ct_pre_off = fd.ct_sparse_off = 0;
            // [ark of round 0][mds][one][r2]: the later round constants are inside the tables
            auto push_fe2 = [&fimg](const Fe& f) {
                for (int k = 0; k < 4; k++) {
                    fimg.push_back((uint32_t)f.v[k]);
                    fimg.push_back((uint32_t)(f.v[k] >> 32));
                }
            };
            fd.ark_off = 0;
            for (uint32_t j = 0; j < t; j++) push_fe2(ctx->ark_mont[j]);
            fd.mds_off = (uint32_t)fimg.size();
            for (const Fe& f : ctx->mds_mont) push_fe2(f);
            fd.one_off = (uint32_t)fimg.size();
            push_fe2(hf_one(*hf));
            fd.r2_off = (uint32_t)fimg.size();
            push_fe2(r2);
            append_fp52_tables(*hf, t, p->full_rounds, p->partial_rounds, ctx->ark_mont, ctx->mds_mont, fd, fimg);
            if (fd.fp_ok) {
                static const bool force_smem = env_int("PSD_FORCE_SMEM", 0) != 0;
                cudaError_t cerr = cudaSuccess;
                ctx->pl_fp = placement_acquire(device, p->field, kFp52PoolKey, fimg, bank_ok && !force_smem, smem_ok, &cerr);
                if (ctx->pl_fp) {
                    if (ctx->pl_fp->kind == kPoolBank) offset_desc(fd, ctx->pl_fp->off);
                    ctx->fdesc = fd;
                } else if (cerr != cudaSuccess) {
                    psd_ctx_destroy(ctx);
                    return fail(PSD_ERR_CUDA, "uploading the variant-5 tables failed: %s", cudaGetErrorString(cerr));
                }
            }
        }
    }

    // Variant 6 (tensor-core linear layer): byte-digit tables in global memory + [ark of round 0][one][r2] staged
    // to shared memory, for the parameter sets tu_tc.cu has kernels for.
    {
        const SpongeKernelEntry* tk = find_kernel(p->field, (int)t, p->alpha, 6, kPoolSmem);
        if (tk && tk->variant == 6 && tk->alpha == p->alpha && env_int("PSD_NO_TC", 0) == 0) {
            TcDesc td{};
            std::vector<uint8_t> tab;
            build_tc_tables(*hf, t, p->full_rounds, p->partial_rounds, ctx->ark_mont, ctx->mds_mont, td, tab);
            if (td.ok) {
                std::vector<uint32_t> small;
                auto push_fe3 = [&small](const Fe& f) {
                    for (int k = 0; k < 4; k++) {
                        small.push_back((uint32_t)f.v[k]);
                        small.push_back((uint32_t)(f.v[k] >> 32));
                    }
                };
                PermDesc sd = base_desc;
                sd.opt_ok = sd.ctab_ok = sd.fp_ok = 0;
                sd.ark_off = 0;
                for (uint32_t j = 0; j < t; j++) push_fe3(ctx->ark_mont[j]);
                sd.mds_off = 0;   // never read by variant 6
                sd.one_off = (uint32_t)small.size();
                push_fe3(hf_one(*hf));
                sd.r2_off = (uint32_t)small.size();
                push_fe3(r2);
                std::vector<uint32_t> tabw(tab.size() / 4);
                memcpy(tabw.data(), tab.data(), tab.size());
                cudaError_t cerr = cudaSuccess;
                ctx->pl_tc = placement_acquire(device, p->field, kTcSmallKey, small, false, true, &cerr);
                if (ctx->pl_tc) ctx->pl_tctab = placement_acquire(device, p->field, kTcTabKey, tabw, false, false, &cerr, true);
                if (ctx->pl_tc && ctx->pl_tctab) {
                    ctx->tdesc = sd;
                    ctx->tcd = td;
                    // variant 7: the same tables in the tcgen05 B-operand layout, where tu_tc5.cu has a kernel
                    const SpongeKernelEntry* k5 = find_kernel(p->field, (int)t, p->alpha, 7, kPoolSmem);
                    if (k5 && k5->variant == 7 && k5->alpha == p->alpha && env_int("PSD_NO_TC5", 0) == 0) {
                        TcDesc td5{};
                        std::vector<uint8_t> tab5;
                        build_tc_tables(*hf, t, p->full_rounds, p->partial_rounds, ctx->ark_mont, ctx->mds_mont, td5, tab5, 1);
                        if (td5.ok) {
                            std::vector<uint32_t> tab5w(tab5.size() / 4);
                            memcpy(tab5w.data(), tab5.data(), tab5.size());
                            ctx->pl_tc5tab = placement_acquire(device, p->field, kTc5TabKey, tab5w, false, false, &cerr, true);
                            if (ctx->pl_tc5tab) ctx->tcd5 = td5;
                        }
                    }
                } else if (cerr != cudaSuccess) {
                    psd_ctx_destroy(ctx);
                    return fail(PSD_ERR_CUDA, "uploading the variant-6 tables failed: %s", cudaGetErrorString(cerr));
                }
            }
        }
    }

    auto cleanup = [&](psd_status s) {
        psd_ctx_destroy(ctx);
        return s;
    };
#define CTX_CUDA(call)                                                                                       \
    do {                                                                                                     \
        cudaError_t e__ = (cudaError_t)(call);                                                               \
        if (e__ != cudaSuccess)                                                                              \
            return cleanup(fail(PSD_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__)));             \
    } while (0)
    cudaDeviceProp prop;
    CTX_CUDA(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    for (int s = 0; s < kStageSlots; s++) CTX_CUDA(cudaStreamCreateWithFlags(&ctx->streams[s], cudaStreamNonBlocking));
    CTX_CUDA(cudaMalloc(&ctx->d_err, sizeof(int32_t)));
    CTX_CUDA(cudaMemset(ctx->d_err, 0, sizeof(int32_t)));
    CTX_CUDA(cudaHostAlloc(&ctx->h_err, sizeof(int32_t), cudaHostAllocDefault));
    *ctx->h_err = 0;
    {
        // [R, R^2] of both fields for the small kernels (convert / synth), once per device
        cudaError_t misc_err = cudaSuccess;
        std::call_once(g_misc_once[device], [&] {
            PoolUploadFn mu = find_uploader(-1, 0);
            std::vector<uint32_t> m;
            for (int fld = 0; fld < 2; fld++) {
                const HostField* h = host_field(fld);
                for (int k = 0; k < 4; k++) { m.push_back((uint32_t)h->one[k]); m.push_back((uint32_t)(h->one[k] >> 32)); }
                for (int k = 0; k < 4; k++) { m.push_back((uint32_t)h->r2[k]); m.push_back((uint32_t)(h->r2[k] >> 32)); }
            }
            if (mu) misc_err = (cudaError_t)mu(0, m.data(), (uint32_t)m.size(), nullptr);
        });
        CTX_CUDA(misc_err);
    }
    ctx->misc_desc = ctx->desc;
    ctx->misc_desc.one_off = misc_one_off(ctx->field);
    ctx->misc_desc.r2_off = misc_r2_off(ctx->field);
    CTX_CUDA(cudaDeviceSynchronize());
#undef CTX_CUDA
    if (!resolve_kernel(ctx).k)
        return cleanup(fail(PSD_ERR_UNSUPPORTED, "no kernel for field %d, t=%u in this build", ctx->field, t));
    *out = ctx;
    return PSD_OK;
}

static void ctx_teardown(psd_ctx* ctx);
static void ctx_release(psd_ctx* ctx) {
    if (ctx->refs.fetch_sub(1) == 1) ctx_teardown(ctx);
}

// The context may outlive this call: batches/sponges created from it hold a reference, so freeing
// them after psd_ctx_destroy is safe (the GPU resources go away with the last of them).
void psd_ctx_destroy(psd_ctx* ctx) {
    if (!ctx) return;
    ctx_release(ctx);
}

static void ctx_teardown(psd_ctx* ctx) {
    {
        DeviceGuard guard(ctx->device);
        for (int s = 0; s < kStageSlots; s++) {
            if (ctx->streams[s]) {
                cudaStreamSynchronize(ctx->streams[s]);
                cudaStreamDestroy(ctx->streams[s]);
            }
            if (ctx->stage_in[s]) cudaFree(ctx->stage_in[s]);
            if (ctx->stage_out[s]) cudaFree(ctx->stage_out[s]);
        }
        if (ctx->d_err) cudaFree(ctx->d_err);
        if (ctx->h_err) cudaFreeHost(ctx->h_err);
        if (ctx->merkle_scratch) cudaFree(ctx->merkle_scratch);
        if (ctx->merkle_leaves) cudaFree(ctx->merkle_leaves);
        if (ctx->multi_roots) cudaFree(ctx->multi_roots);
        if (ctx->multi_event) cudaEventDestroy(ctx->multi_event);
        placement_release(ctx->pl_int);
        placement_release(ctx->pl_fp);
        placement_release(ctx->pl_tc);
        placement_release(ctx->pl_tctab);
        placement_release(ctx->pl_tc5tab);
    }
    delete ctx;
}

psd_status psd_ctx_get_params(const psd_ctx* ctx, psd_params* out) {
    if (!ctx || !out) return fail(PSD_ERR_ARG, "null argument");
    out->field = ctx->field;
    out->full_rounds = ctx->full_rounds;
    out->partial_rounds = ctx->partial_rounds;
    out->alpha = ctx->alpha;
    out->rate = ctx->rate;
    out->capacity = ctx->capacity;
    out->ark = nullptr;
    out->mds = nullptr;
    out->repr = PSD_REPR_MONTGOMERY;
    return PSD_OK;
}

psd_status psd_ctx_set_variant(psd_ctx* ctx, int variant) {
    if (!ctx) return fail(PSD_ERR_ARG, "null context");
    if (variant < 0 || variant > 7) return fail(PSD_ERR_ARG, "variant must be 0..7");
    ctx->variant = variant;
    return PSD_OK;
}

psd_status psd_ctx_active_variant(const psd_ctx* ctx, int* variant) {
    if (!ctx || !variant) return fail(PSD_ERR_ARG, "null argument");
    const SpongeKernelEntry* k = resolve_kernel(ctx).k;
    if (!k) return fail(PSD_ERR_UNSUPPORTED, "no kernel for field %d, t=%u", ctx->field, ctx->t);
    *variant = k->variant;
    return PSD_OK;
}

psd_status psd_ctx_describe(const psd_ctx* ctx, psd_ctx_info* out) {
    if (!ctx || !out) return fail(PSD_ERR_ARG, "null argument");
    const Resolved r = resolve_kernel(ctx);
    if (!r.k) return fail(PSD_ERR_UNSUPPORTED, "no kernel for field %d, t=%u", ctx->field, ctx->t);
    memset(out, 0, sizeof *out);
    out->variant = r.k->variant;
    out->constants_home = r.pl->kind == kPoolBank ? PSD_CONSTANTS_BANK : PSD_CONSTANTS_SHARED;
    out->image_bytes = (uint32_t)(r.pl->words.size() * 4);
    out->image_shared_by = (uint32_t)r.pl->refs;
    snprintf(out->kernel_name, sizeof out->kernel_name, "%s", r.k->name);
    return PSD_OK;
}

// ------------------------------------------------------------------------------------------------
// Memory helpers
// ------------------------------------------------------------------------------------------------
psd_status psd_host_alloc(size_t bytes, void** out) {
    if (!out) return fail(PSD_ERR_ARG, "null output");
    *out = nullptr;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? PSD_ERR_NOMEM : PSD_ERR_CUDA, "cudaHostAlloc(%zu): %s", bytes,
                    cudaGetErrorString(e));
    }
    return PSD_OK;
}
void psd_host_free(void* p) {
    if (p) cudaFreeHost(p);
}
psd_status psd_device_alloc(psd_ctx* ctx, size_t bytes, void** out) {
    if (!ctx || !out) return fail(PSD_ERR_ARG, "null argument");
    *out = nullptr;
    DeviceGuard guard(ctx->device);
    cudaError_t e = cudaMalloc(out, bytes ? bytes : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? PSD_ERR_NOMEM : PSD_ERR_CUDA, "cudaMalloc(%zu): %s", bytes,
                    cudaGetErrorString(e));
    }
    return PSD_OK;
}
void psd_device_free(psd_ctx* ctx, void* p) {
    if (!ctx || !p) return;
    DeviceGuard guard(ctx->device);
    cudaFree(p);
}
psd_status psd_memcpy_h2d(psd_ctx* ctx, void* dst, const void* src, size_t bytes, void* stream) {
    if (!ctx || !dst || !src) return fail(PSD_ERR_ARG, "null argument");
    DeviceGuard guard(ctx->device);
    PSD_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return PSD_OK;
}
psd_status psd_memcpy_d2h(psd_ctx* ctx, void* dst, const void* src, size_t bytes, void* stream) {
    if (!ctx || !dst || !src) return fail(PSD_ERR_ARG, "null argument");
    DeviceGuard guard(ctx->device);
    PSD_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return PSD_OK;
}
psd_status psd_stream_synchronize(psd_ctx* ctx, void* stream) {
    if (!ctx) return fail(PSD_ERR_ARG, "null context");
    DeviceGuard guard(ctx->device);
    PSD_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return check_err_flag(ctx, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// Representation change, synthetic inputs, K0
// ------------------------------------------------------------------------------------------------
static psd_status convert(psd_ctx* ctx, uint64_t* elems, size_t n, int mem, void* stream, int to_m) {
    if (!ctx || (!elems && n)) return fail(PSD_ERR_ARG, "null argument");
    if (!valid_enum(mem, 1)) return fail(PSD_ERR_ARG, "bad mem");
    if (n == 0) return PSD_OK;
    DeviceGuard guard(ctx->device);
    if (mem == PSD_MEM_DEVICE) {
        PSD_CUDA(launch_convert(ctx->field, ctx->misc_desc, as_u32(elems), n, to_m, ctx->d_err, misc_cfg(ctx, n), stream));
        g_launches++;
        return PSD_OK;
    }
    std::lock_guard<std::mutex> lk(ctx->stage_mu);
    const uint64_t chunk = std::min<uint64_t>(n, 1ull << 22);
    psd_status st = ensure_staging(ctx, chunk * 32, 0);
    if (st != PSD_OK) return st;
    int slot = 0;
    for (uint64_t first = 0; first < n; first += chunk, slot = (slot + 1) % kStageSlots) {
        const uint64_t cnt = std::min<uint64_t>(chunk, n - first);
        cudaStream_t s = ctx->streams[slot];
        PSD_CUDA(cudaMemcpyAsync(ctx->stage_in[slot], elems + first * 4, cnt * 32, cudaMemcpyHostToDevice, s));
        PSD_CUDA(launch_convert(ctx->field, ctx->misc_desc, (uint32_t*)ctx->stage_in[slot], cnt, to_m, ctx->d_err,
                                misc_cfg(ctx, cnt), s));
        g_launches++;
        PSD_CUDA(cudaMemcpyAsync(elems + first * 4, ctx->stage_in[slot], cnt * 32, cudaMemcpyDeviceToHost, s));
    }
    for (int s = 0; s < kStageSlots; s++) PSD_CUDA(cudaStreamSynchronize(ctx->streams[s]));
    return check_err_flag(ctx, ctx->streams[0]);
}
psd_status psd_to_montgomery(psd_ctx* ctx, uint64_t* elems, size_t n, int mem, void* stream) {
    return convert(ctx, elems, n, mem, stream, 1);
}
psd_status psd_from_montgomery(psd_ctx* ctx, uint64_t* elems, size_t n, int mem, void* stream) {
    return convert(ctx, elems, n, mem, stream, 0);
}

psd_status psd_synth_fill(psd_ctx* ctx, uint64_t* dev_elems, size_t n, uint64_t seed, uint64_t first_index, int repr,
                          void* stream) {
    if (!ctx || (!dev_elems && n)) return fail(PSD_ERR_ARG, "null argument");
    if (!valid_enum(repr, 1)) return fail(PSD_ERR_ARG, "bad repr");
    if (n == 0) return PSD_OK;
    DeviceGuard guard(ctx->device);
    PSD_CUDA(launch_synth(ctx->field, ctx->misc_desc, as_u32(dev_elems), n, seed, first_index, repr, misc_cfg(ctx, n), stream));
    g_launches++;
    return PSD_OK;
}

psd_status psd_imad_peak(psd_ctx* ctx, int kind, double* per_second, double* elapsed_ms) {
    if (!ctx || !per_second) return fail(PSD_ERR_ARG, "null argument");
    if (kind < 0 || kind > 3) return fail(PSD_ERR_ARG, "kind must be 0..3");
    DeviceGuard guard(ctx->device);
    uint32_t* scratch = nullptr;
    PSD_CUDA(cudaMalloc(&scratch, 1024 * 4));
    std::vector<uint32_t> init(1024);
    for (int i = 0; i < 1024; i++) init[i] = 0x9e3779b9u * (i + 1) + 12345u;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    psd_status result = PSD_OK;
    float ms = 0;
    LaunchCfg cfg;
    cfg.block = kind >= 2 ? 128 : 256;
    cfg.grid = ctx->sm_count * 8;
    const uint32_t iters = kind >= 2 ? 4096 : 16384;
    do {
#define K0_CUDA(call)                                                                           \
    if ((cudaError_t)(call) != cudaSuccess) {                                                   \
        result = fail(PSD_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(cudaGetLastError())); \
        break;                                                                                  \
    }
        K0_CUDA(cudaMemcpy(scratch, init.data(), 4096, cudaMemcpyHostToDevice));
        K0_CUDA(cudaEventCreate(&e0));
        K0_CUDA(cudaEventCreate(&e1));
        cudaStream_t s = ctx->streams[0];
        for (int w = 0; w < 2; w++) K0_CUDA(launch_imad_peak(ctx->field, kind, scratch, iters, cfg, s));
        K0_CUDA(cudaEventRecord(e0, s));
        const int reps = 5;
        for (int r = 0; r < reps; r++) K0_CUDA(launch_imad_peak(ctx->field, kind, scratch, iters, cfg, s));
        K0_CUDA(cudaEventRecord(e1, s));
        K0_CUDA(cudaEventSynchronize(e1));
        K0_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        g_launches += 7;
        const double threads = (double)cfg.grid * cfg.block;
        const double per_thread = kind >= 2 ? 2.0 * iters : 64.0 * iters;
        *per_second = threads * per_thread * reps / (ms * 1e-3);
        if (elapsed_ms) *elapsed_ms = ms / reps;
#undef K0_CUDA
    } while (0);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(scratch);
    return result;
}

// ------------------------------------------------------------------------------------------------
// Batch entry points
// ------------------------------------------------------------------------------------------------
psd_status psd_permute_batch(psd_ctx* ctx, uint64_t* states, size_t n, int layout, int repr, int mem, void* stream) {
    psd_status st = check_batch_args(ctx, n ? (void*)states : (void*)ctx, layout, repr, mem);
    if (st != PSD_OK) return st;
    if (n == 0) return PSD_OK;
    DeviceGuard guard(ctx->device);
    auto fill = [&](SpongeArgs& a, void* d_in, void* /*d_out*/) {
        a.state_in = (const uint32_t*)d_in;
        a.state_in_layout = layout;
        a.state_in_repr = repr;
        a.state_out = (uint32_t*)d_in;
        a.state_out_layout = layout;
        a.state_out_repr = repr;
        a.n_raw = 1;
    };
    if (mem == PSD_MEM_DEVICE) {
        SpongeArgs a = base_args(ctx, n);
        fill(a, states, states);
        return launch(ctx, a, (cudaStream_t)stream);
    }
    StagedCall c{n, states, ctx->t, states, ctx->t, 0, layout, true};
    return run_staged(ctx, c, fill);
}

static psd_status absorb_squeeze(psd_ctx* ctx, const uint64_t* inputs, size_t n, uint32_t n_absorb, uint32_t n_squeeze,
                                 uint64_t* outputs, uint8_t* out_bytes, uint32_t n_bytes, int layout, int repr, int mem,
                                 void* stream) {
    psd_status st = check_batch_args(ctx, (n && n_absorb) ? (const void*)inputs : (const void*)ctx, layout, repr, mem);
    if (st != PSD_OK) return st;
    const bool bytes_mode = out_bytes != nullptr || (outputs == nullptr && n_bytes);
    if (n && ((bytes_mode && n_bytes && !out_bytes) || (!bytes_mode && n_squeeze && !outputs)))
        return fail(PSD_ERR_ARG, "null output buffer");
    if (n == 0) return PSD_OK;
    if (bytes_mode) {
        const uint32_t usable = (uint32_t)(ctx->hf->bits - 1) / 8;
        n_squeeze = (n_bytes + usable - 1) / usable;   // mod.rs:259
    }
    DeviceGuard guard(ctx->device);
    auto fill = [&](SpongeArgs& a, void* d_in, void* d_out) {
        a.in = (const uint32_t*)d_in;
        a.n_absorb = n_absorb;
        a.in_layout = layout;
        a.in_repr = repr;
        a.do_squeeze = 1;
        a.n_squeeze = n_squeeze;
        if (bytes_mode) {
            a.out_bytes = (uint8_t*)d_out;
            a.n_bytes = n_bytes;
        } else {
            a.out = (uint32_t*)d_out;
            a.out_layout = layout;
            a.out_repr = repr;
        }
    };
    if (mem == PSD_MEM_DEVICE) {
        SpongeArgs a = base_args(ctx, n);
        fill(a, (void*)inputs, bytes_mode ? (void*)out_bytes : (void*)outputs);
        return launch(ctx, a, (cudaStream_t)stream);
    }
    StagedCall c{n, inputs, n_absorb, bytes_mode ? (void*)out_bytes : (void*)outputs, bytes_mode ? 0 : n_squeeze,
                 bytes_mode ? n_bytes : 0, layout, false};
    if (bytes_mode && n_bytes == 0) c.out_bytes_per = 0, c.out_w = 0;
    return run_staged(ctx, c, fill);
}

psd_status psd_absorb_squeeze_batch(psd_ctx* ctx, const uint64_t* inputs, size_t n, uint32_t n_absorb, uint32_t n_squeeze,
                                    uint64_t* outputs, int layout, int repr, int mem, void* stream) {
    return absorb_squeeze(ctx, inputs, n, n_absorb, n_squeeze, outputs, nullptr, 0, layout, repr, mem, stream);
}

psd_status psd_absorb_squeeze_bytes_batch(psd_ctx* ctx, const uint64_t* inputs, size_t n, uint32_t n_absorb,
                                          uint32_t n_bytes, uint8_t* outputs, int layout, int repr, int mem, void* stream) {
    if (n && n_bytes && !outputs) return fail(PSD_ERR_ARG, "null output buffer");
    static uint8_t dummy;
    return absorb_squeeze(ctx, inputs, n, n_absorb, 0, nullptr, outputs ? outputs : &dummy, n_bytes, layout, repr, mem, stream);
}

psd_status psd_compress_pairs(psd_ctx* ctx, const uint64_t* in, size_t n_pairs, uint64_t* out, int repr, int mem,
                              void* stream) {
    if (ctx && ctx->rate < 2) return fail(PSD_ERR_UNSUPPORTED, "2-to-1 compression needs rate >= 2");
    return absorb_squeeze(ctx, in, n_pairs, 2, 1, out, nullptr, 0, PSD_LAYOUT_AOS, repr, mem, stream);
}

psd_status psd_merkle_root(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, uint64_t* root, int repr, int mem,
                           void* stream) {
    psd_status st = check_batch_args(ctx, leaves, PSD_LAYOUT_AOS, repr, mem);
    if (st != PSD_OK) return st;
    if (!root) return fail(PSD_ERR_ARG, "null root");
    if (ctx->rate < 2) return fail(PSD_ERR_UNSUPPORTED, "2-to-1 compression needs rate >= 2");
    if (n_leaves == 0 || (n_leaves & (n_leaves - 1))) return fail(PSD_ERR_ARG, "n_leaves must be a power of two");
    DeviceGuard guard(ctx->device);
    cudaStream_t s = mem == PSD_MEM_DEVICE ? (cudaStream_t)stream : ctx->streams[0];
    if (n_leaves == 1) {
        const cudaMemcpyKind kind = mem == PSD_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToHost;
        PSD_CUDA(cudaMemcpyAsync(root, leaves, 32, kind, s));
        PSD_CUDA(cudaStreamSynchronize(s));
        return PSD_OK;
    }
    std::lock_guard<std::mutex> lk(ctx->merkle_mu);
    psd_status result = PSD_OK;
    do {
        const uint32_t* src = as_u32(leaves);
        if (mem == PSD_MEM_HOST) {
            result = merkle_upload_leaves(ctx, leaves, n_leaves, s, &src);
            if (result != PSD_OK) break;
        }
        const uint32_t* d_root = nullptr;
        result = merkle_reduce_async(ctx, src, repr, n_leaves, mem == PSD_MEM_DEVICE ? as_u32(root) : nullptr, repr, s, &d_root);
        if (result != PSD_OK) break;
        if (mem == PSD_MEM_HOST && cudaMemcpyAsync(root, d_root, 32, cudaMemcpyDeviceToHost, s) != cudaSuccess) {
            result = fail(PSD_ERR_CUDA, "root download failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        if (cudaStreamSynchronize(s) != cudaSuccess) {
            result = fail(PSD_ERR_CUDA, "Merkle levels failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        result = check_err_flag(ctx, s);
    } while (0);
    cudaStreamSynchronize(s);
    return result;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Lock-step batches
// ------------------------------------------------------------------------------------------------
struct psd_batch {
    psd_ctx* ctx = nullptr;
    size_t n = 0;
    uint32_t* d_state = nullptr;   // SoA, Montgomery, n * t elements
    int mode = kModeAbsorbing;
    uint32_t index = 0;
    cudaStream_t stream = nullptr;      // stream in use (own_stream unless the caller set one)
    cudaStream_t own_stream = nullptr;
    void* d_io = nullptr;          // device staging for host-side inputs/outputs
    size_t d_io_bytes = 0;
    // this batch's own non-canonical-input flag, so that batches of one context used from different
    // threads cannot steal or clear each other's
    int32_t* d_err = nullptr;      // lives in the tail of the d_state allocation
    int32_t h_err = 0;
};

namespace {

psd_status batch_io(psd_batch* b, size_t bytes) {
    if (bytes <= b->d_io_bytes) return PSD_OK;
    if (b->d_io) cudaFree(b->d_io);
    b->d_io = nullptr;
    b->d_io_bytes = 0;
    if (cudaMalloc(&b->d_io, bytes) != cudaSuccess) {
        cudaGetLastError();
        return fail(PSD_ERR_NOMEM, "cudaMalloc(%zu) failed", bytes);
    }
    b->d_io_bytes = bytes;
    return PSD_OK;
}

void batch_state_args(const psd_batch* b, SpongeArgs& a) {
    a.err = b->d_err;
    a.state_in = b->d_state;
    a.state_in_layout = kLayoutSoa;
    a.state_in_repr = kReprMontgomery;
    a.state_out = b->d_state;
    a.state_out_layout = kLayoutSoa;
    a.state_out_repr = kReprMontgomery;
    a.mode = b->mode;
    a.index = b->index;
}

}  // namespace

extern "C" {

psd_status psd_batch_new(psd_ctx* ctx, size_t n, psd_batch** out) {
    if (!ctx || !out) return fail(PSD_ERR_ARG, "null argument");
    *out = nullptr;
    if (n == 0) return fail(PSD_ERR_ARG, "empty batch");
    DeviceGuard guard(ctx->device);
    psd_batch* b = new (std::nothrow) psd_batch();
    if (!b) return fail(PSD_ERR_NOMEM, "out of host memory");
    b->ctx = ctx;
    ctx->refs.fetch_add(1);
    b->n = n;
    const size_t bytes = n * ctx->t * 32;
    if (cudaMalloc(&b->d_state, bytes + 16) != cudaSuccess) {
        cudaGetLastError();
        ctx_release(ctx);
        delete b;
        return fail(PSD_ERR_NOMEM, "cudaMalloc(%zu) for batch state failed", bytes);
    }
    if (cudaStreamCreateWithFlags(&b->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        (b->stream = b->own_stream, false) ||
        (b->d_err = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(b->d_state) + bytes), false) ||
        cudaMemsetAsync(b->d_state, 0, bytes + 16, b->stream) != cudaSuccess ||   // vec![F::zero(); t], mod.rs:220
        cudaStreamSynchronize(b->stream) != cudaSuccess) {
        psd_status st = fail(PSD_ERR_CUDA, "batch init failed: %s", cudaGetErrorString(cudaGetLastError()));
        psd_batch_free(b);
        return st;
    }
    *out = b;
    return PSD_OK;
}

psd_status psd_batch_clone(const psd_batch* b, psd_batch** out) {
    if (!b || !out) return fail(PSD_ERR_ARG, "null argument");
    psd_status st = psd_batch_new(b->ctx, b->n, out);
    if (st != PSD_OK) return st;
    DeviceGuard guard(b->ctx->device);
    psd_batch* c = *out;
    c->mode = b->mode;
    c->index = b->index;
    if (cudaStreamSynchronize(b->stream) != cudaSuccess ||
        cudaMemcpyAsync(c->d_state, b->d_state, b->n * b->ctx->t * 32, cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
        st = fail(PSD_ERR_CUDA, "batch clone failed: %s", cudaGetErrorString(cudaGetLastError()));
        psd_batch_free(c);
        *out = nullptr;
        return st;
    }
    return PSD_OK;
}

void psd_batch_free(psd_batch* b) {
    if (!b) return;
    {
        DeviceGuard guard(b->ctx->device);
        if (b->stream) cudaStreamSynchronize(b->stream);
        if (b->own_stream) cudaStreamDestroy(b->own_stream);
        if (b->d_state) cudaFree(b->d_state);
        if (b->d_io) cudaFree(b->d_io);
    }
    ctx_release(b->ctx);
    delete b;
}

size_t psd_batch_size(const psd_batch* b) { return b ? b->n : 0; }

psd_status psd_batch_set_stream(psd_batch* b, void* stream) {
    if (!b) return fail(PSD_ERR_ARG, "null batch");
    DeviceGuard guard(b->ctx->device);
    PSD_CUDA(cudaStreamSynchronize(b->stream));
    b->stream = stream ? (cudaStream_t)stream : b->own_stream;
    return PSD_OK;
}

psd_status psd_batch_absorb(psd_batch* b, const uint64_t* elems, uint32_t k, int layout, int repr, int mem) {
    if (!b) return fail(PSD_ERR_ARG, "null batch");
    if (k == 0) return PSD_OK;   // empty input: no-op, mod.rs:234-236
    psd_status st = check_batch_args(b->ctx, elems, layout, repr, mem);
    if (st != PSD_OK) return st;
    psd_ctx* ctx = b->ctx;
    DeviceGuard guard(ctx->device);
    const uint32_t* d_in = as_u32(elems);
    if (mem == PSD_MEM_HOST) {
        const size_t bytes = b->n * (size_t)k * 32;
        st = batch_io(b, bytes);
        if (st != PSD_OK) return st;
        PSD_CUDA(cudaMemcpyAsync(b->d_io, elems, bytes, cudaMemcpyHostToDevice, b->stream));
        d_in = (const uint32_t*)b->d_io;
    }
    SpongeArgs a = base_args(ctx, b->n);
    batch_state_args(b, a);
    a.in = d_in;
    a.n_absorb = k;
    a.in_layout = layout;
    a.in_repr = repr;
    st = launch(ctx, a, b->stream);
    if (st != PSD_OK) return st;
    PSD_CUDA(cudaStreamSynchronize(b->stream));
    if (repr == PSD_REPR_CANONICAL) {
        // A rejected input leaves mode/index where they were (the reference would have panicked inside
        // from_bigint before touching the sponge); the states of the batch are then unspecified and must
        // be restored with psd_batch_set_state before it is used again.
        st = read_err_flag(&b->h_err, b->d_err, b->stream);
        if (st != PSD_OK) return st;
    }
    advance_mode(ctx->rate, k, false, 0, &b->mode, &b->index);
    return PSD_OK;
}

static psd_status batch_squeeze(psd_batch* b, uint32_t k, uint64_t* out, uint8_t* out_bytes, uint32_t n_bytes, int layout,
                                int repr, int mem) {
    psd_ctx* ctx = b->ctx;
    DeviceGuard guard(ctx->device);
    const bool bytes_mode = out_bytes != nullptr;
    const size_t out_size = bytes_mode ? b->n * (size_t)n_bytes : b->n * (size_t)k * 32;
    void* d_out = bytes_mode ? (void*)out_bytes : (void*)out;
    if (mem == PSD_MEM_HOST) {
        psd_status st = batch_io(b, std::max<size_t>(out_size, 16));
        if (st != PSD_OK) return st;
        d_out = b->d_io;
    }
    SpongeArgs a = base_args(ctx, b->n);
    batch_state_args(b, a);
    a.do_squeeze = 1;
    a.n_squeeze = k;
    if (bytes_mode) {
        a.out_bytes = (uint8_t*)d_out;
        a.n_bytes = n_bytes;
    } else {
        a.out = (uint32_t*)d_out;
        a.out_layout = layout;
        a.out_repr = repr;
    }
    psd_status st = launch(ctx, a, b->stream);
    if (st != PSD_OK) return st;
    advance_mode(ctx->rate, 0, true, k, &b->mode, &b->index);
    if (mem == PSD_MEM_HOST && out_size)
        PSD_CUDA(cudaMemcpyAsync(bytes_mode ? (void*)out_bytes : (void*)out, d_out, out_size, cudaMemcpyDeviceToHost, b->stream));
    PSD_CUDA(cudaStreamSynchronize(b->stream));
    return PSD_OK;
}

psd_status psd_batch_squeeze_native(psd_batch* b, uint32_t k, uint64_t* out, int layout, int repr, int mem) {
    if (!b) return fail(PSD_ERR_ARG, "null batch");
    static uint64_t dummy[4];
    psd_status st = check_batch_args(b->ctx, k ? (void*)out : (void*)dummy, layout, repr, mem);
    if (st != PSD_OK) return st;
    return batch_squeeze(b, k, out ? out : dummy, nullptr, 0, layout, repr, k ? mem : PSD_MEM_HOST);
}

psd_status psd_batch_squeeze_bytes(psd_batch* b, uint32_t n_bytes, uint8_t* out, int mem) {
    if (!b) return fail(PSD_ERR_ARG, "null batch");
    if (!valid_enum(mem, 1)) return fail(PSD_ERR_ARG, "bad mem");
    if (n_bytes && !out) return fail(PSD_ERR_ARG, "null output");
    static uint8_t dummy;
    const uint32_t usable = (uint32_t)(b->ctx->hf->bits - 1) / 8;
    const uint32_t k = (n_bytes + usable - 1) / usable;
    return batch_squeeze(b, k, nullptr, out ? out : &dummy, n_bytes, PSD_LAYOUT_AOS, PSD_REPR_CANONICAL,
                         n_bytes ? mem : PSD_MEM_HOST);
}

// SpongeExt::into_state / from_state (lib.rs:188-195): a pure layout/representation change done by
// the sponge kernel with no absorb, squeeze or permutation in between load and store.
psd_status psd_batch_get_state(const psd_batch* b, uint64_t* states, int layout, int repr, int mem, int* mode,
                               uint32_t* index) {
    if (!b) return fail(PSD_ERR_ARG, "null batch");
    psd_status st = check_batch_args(b->ctx, states, layout, repr, mem);
    if (st != PSD_OK) return st;
    psd_ctx* ctx = b->ctx;
    DeviceGuard guard(ctx->device);
    psd_batch* bm = const_cast<psd_batch*>(b);
    const size_t bytes = b->n * ctx->t * 32;
    uint32_t* d_user = as_u32(states);
    if (mem == PSD_MEM_HOST) {
        st = batch_io(bm, bytes);
        if (st != PSD_OK) return st;
        d_user = (uint32_t*)bm->d_io;
    }
    SpongeArgs a = base_args(ctx, b->n);
    a.state_in = b->d_state;
    a.state_in_layout = kLayoutSoa;
    a.state_in_repr = kReprMontgomery;
    a.state_out = d_user;
    a.state_out_layout = layout;
    a.state_out_repr = repr;
    st = launch(ctx, a, b->stream);
    if (st != PSD_OK) return st;
    if (mem == PSD_MEM_HOST) PSD_CUDA(cudaMemcpyAsync(states, d_user, bytes, cudaMemcpyDeviceToHost, b->stream));
    PSD_CUDA(cudaStreamSynchronize(b->stream));
    if (mode) *mode = b->mode;
    if (index) *index = b->index;
    return PSD_OK;
}

psd_status psd_batch_set_state(psd_batch* b, const uint64_t* states, int layout, int repr, int mem, int mode,
                               uint32_t index) {
    if (!b) return fail(PSD_ERR_ARG, "null batch");
    psd_status st = check_batch_args(b->ctx, states, layout, repr, mem);
    if (st != PSD_OK) return st;
    psd_ctx* ctx = b->ctx;
    if (!valid_enum(mode, 1) || index > ctx->rate) return fail(PSD_ERR_ARG, "bad mode/index");
    DeviceGuard guard(ctx->device);
    const size_t bytes = b->n * ctx->t * 32;
    const uint32_t* d_user = as_u32(states);
    if (mem == PSD_MEM_HOST) {
        st = batch_io(b, bytes);
        if (st != PSD_OK) return st;
        PSD_CUDA(cudaMemcpyAsync(b->d_io, states, bytes, cudaMemcpyHostToDevice, b->stream));
        d_user = (const uint32_t*)b->d_io;
    }
    SpongeArgs a = base_args(ctx, b->n);
    a.state_in = d_user;
    a.state_in_layout = layout;
    a.state_in_repr = repr;
    a.state_out = b->d_state;
    a.state_out_layout = kLayoutSoa;
    a.state_out_repr = kReprMontgomery;
    a.err = b->d_err;
    st = launch(ctx, a, b->stream);
    if (st != PSD_OK) return st;
    PSD_CUDA(cudaStreamSynchronize(b->stream));
    if (repr == PSD_REPR_CANONICAL) {
        st = read_err_flag(&b->h_err, b->d_err, b->stream);
        if (st != PSD_OK) return st;
    }
    b->mode = mode;
    b->index = index;
    return PSD_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Single-sponge drop-in: a batch of one.
// ------------------------------------------------------------------------------------------------
struct psd_sponge {
    psd_batch* b = nullptr;
};

extern "C" {

psd_status psd_sponge_new(psd_ctx* ctx, psd_sponge** out) {
    if (!ctx || !out) return fail(PSD_ERR_ARG, "null argument");
    *out = nullptr;
    psd_sponge* s = new (std::nothrow) psd_sponge();
    if (!s) return fail(PSD_ERR_NOMEM, "out of host memory");
    psd_status st = psd_batch_new(ctx, 1, &s->b);
    if (st != PSD_OK) {
        delete s;
        return st;
    }
    *out = s;
    return PSD_OK;
}

psd_status psd_sponge_clone(const psd_sponge* s, psd_sponge** out) {
    if (!s || !out) return fail(PSD_ERR_ARG, "null argument");
    *out = nullptr;
    psd_sponge* c = new (std::nothrow) psd_sponge();
    if (!c) return fail(PSD_ERR_NOMEM, "out of host memory");
    psd_status st = psd_batch_clone(s->b, &c->b);
    if (st != PSD_OK) {
        delete c;
        return st;
    }
    *out = c;
    return PSD_OK;
}

void psd_sponge_free(psd_sponge* s) {
    if (!s) return;
    psd_batch_free(s->b);
    delete s;
}

psd_status psd_sponge_absorb(psd_sponge* s, const uint64_t* elems, size_t n, int repr) {
    if (!s) return fail(PSD_ERR_ARG, "null sponge");
    if (n == 0) return PSD_OK;
    if (n > 0xffffffffull) return fail(PSD_ERR_ARG, "too many elements");
    return psd_batch_absorb(s->b, elems, (uint32_t)n, PSD_LAYOUT_AOS, repr, PSD_MEM_HOST);
}

psd_status psd_sponge_absorb_bytes(psd_sponge* s, const uint8_t* bytes, size_t n) {
    if (!s || (!bytes && n)) return fail(PSD_ERR_ARG, "null argument");
    // absorb.rs:135-139: (len as u64 LE) || bytes, cut into (MODULUS_BIT_SIZE-1)/8-byte LE chunks
    const size_t usable = (size_t)(s->b->ctx->hf->bits - 1) / 8;
    std::vector<uint8_t> buf(8 + n);
    const uint64_t len = (uint64_t)n;
    for (int i = 0; i < 8; i++) buf[i] = (uint8_t)(len >> (8 * i));
    if (n) memcpy(buf.data() + 8, bytes, n);
    const size_t n_el = (buf.size() + usable - 1) / usable;
    std::vector<uint64_t> elems(n_el * 4, 0);
    for (size_t i = 0; i < buf.size(); i++) {
        const size_t e = i / usable, off = i % usable;
        elems[e * 4 + off / 8] |= (uint64_t)buf[i] << (8 * (off % 8));
    }
    return psd_sponge_absorb(s, elems.data(), n_el, PSD_REPR_CANONICAL);
}

psd_status psd_sponge_squeeze_native(psd_sponge* s, size_t n, uint64_t* out, int repr) {
    if (!s) return fail(PSD_ERR_ARG, "null sponge");
    if (n > 0xffffffffull) return fail(PSD_ERR_ARG, "too many elements");
    return psd_batch_squeeze_native(s->b, (uint32_t)n, out, PSD_LAYOUT_AOS, repr, PSD_MEM_HOST);
}

psd_status psd_sponge_squeeze_bytes(psd_sponge* s, size_t n_bytes, uint8_t* out) {
    if (!s) return fail(PSD_ERR_ARG, "null sponge");
    if (n_bytes > 0xffffffffull) return fail(PSD_ERR_ARG, "too many bytes");
    return psd_batch_squeeze_bytes(s->b, (uint32_t)n_bytes, out, PSD_MEM_HOST);
}

psd_status psd_sponge_squeeze_bits(psd_sponge* s, size_t n_bits, uint8_t* out) {
    if (!s || (!out && n_bits)) return fail(PSD_ERR_ARG, "null argument");
    // mod.rs:272-286: MODULUS_BIT_SIZE-1 LE bits of each canonical element
    const size_t usable = (size_t)s->b->ctx->hf->bits - 1;
    const size_t n_el = (n_bits + usable - 1) / usable;
    std::vector<uint64_t> elems(std::max<size_t>(n_el, 1) * 4);
    psd_status st = psd_sponge_squeeze_native(s, n_el, elems.data(), PSD_REPR_CANONICAL);
    if (st != PSD_OK) return st;
    for (size_t i = 0; i < n_bits; i++) {
        const size_t e = i / usable, bit = i % usable;
        out[i] = (uint8_t)((elems[e * 4 + bit / 64] >> (bit % 64)) & 1);
    }
    return PSD_OK;
}

psd_status psd_sponge_squeeze_field_elements(psd_sponge* s, int32_t target_field, size_t n, uint64_t* out) {
    if (!s || (!out && n)) return fail(PSD_ERR_ARG, "null argument");
    const HostField* f2 = host_field(target_field);
    if (!f2) return fail(PSD_ERR_ARG, "unknown target field %d", target_field);
    if (target_field == s->b->ctx->field)   // TypeId equal / same characteristic: native, mod.rs:288-317
        return psd_sponge_squeeze_native(s, n, out, PSD_REPR_CANONICAL);
    if (n == 0) return PSD_OK;   // lib.rs:65-67
    // lib.rs:61-100: (MODULUS_BIT_SIZE_F2 - 1) bits per element, LE-packed, from_le_bytes_mod_order
    const size_t nb = (size_t)f2->bits - 1;
    std::vector<uint8_t> bits(nb * n);
    psd_status st = psd_sponge_squeeze_bits(s, nb * n, bits.data());
    if (st != PSD_OK) return st;
    for (size_t k = 0; k < n; k++) {
        Fe v = hf_zero();
        for (size_t i = 0; i < nb; i++) v.v[i / 64] |= (uint64_t)bits[k * nb + i] << (i % 64);
        v = hf_reduce(*f2, v);
        memcpy(out + 4 * k, v.v, 32);
    }
    return PSD_OK;
}

psd_status psd_sponge_fork(const psd_sponge* s, const uint8_t* domain, size_t n, psd_sponge** out) {
    if (!s || !out || (!domain && n)) return fail(PSD_ERR_ARG, "null argument");
    // lib.rs:149-157: clone; absorb(&(len as u64 LE bytes || domain) as Vec<u8>)
    psd_status st = psd_sponge_clone(s, out);
    if (st != PSD_OK) return st;
    std::vector<uint8_t> input(8 + n);
    const uint64_t len = (uint64_t)n;
    for (int i = 0; i < 8; i++) input[i] = (uint8_t)(len >> (8 * i));
    if (n) memcpy(input.data() + 8, domain, n);
    st = psd_sponge_absorb_bytes(*out, input.data(), input.size());
    if (st != PSD_OK) {
        psd_sponge_free(*out);
        *out = nullptr;
    }
    return st;
}

psd_status psd_sponge_get_state(const psd_sponge* s, uint64_t* state, int repr, int* mode, uint32_t* index) {
    if (!s) return fail(PSD_ERR_ARG, "null sponge");
    return psd_batch_get_state(s->b, state, PSD_LAYOUT_AOS, repr, PSD_MEM_HOST, mode, index);
}

psd_status psd_sponge_set_state(psd_sponge* s, const uint64_t* state, int repr, int mode, uint32_t index) {
    if (!s) return fail(PSD_ERR_ARG, "null sponge");
    return psd_batch_set_state(s->b, state, PSD_LAYOUT_AOS, repr, PSD_MEM_HOST, mode, index);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Full Merkle tree (all internal nodes) and authentication paths
// ------------------------------------------------------------------------------------------------
extern "C" psd_status psd_merkle_tree(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, uint64_t* nodes, int repr,
                                      int mem, void* stream) {
    psd_status st = check_batch_args(ctx, leaves, PSD_LAYOUT_AOS, repr, mem);
    if (st != PSD_OK) return st;
    if (!nodes) return fail(PSD_ERR_ARG, "null nodes buffer");
    if (ctx->rate < 2) return fail(PSD_ERR_UNSUPPORTED, "2-to-1 compression needs rate >= 2");
    if (n_leaves < 2 || (n_leaves & (n_leaves - 1))) return fail(PSD_ERR_ARG, "n_leaves must be a power of two >= 2");
    DeviceGuard guard(ctx->device);
    cudaStream_t s = mem == PSD_MEM_DEVICE ? (cudaStream_t)stream : ctx->streams[0];
    const uint32_t* d_leaves = as_u32(leaves);
    uint32_t* d_nodes = as_u32(nodes);
    std::unique_lock<std::mutex> lk(ctx->merkle_mu, std::defer_lock);
    if (mem == PSD_MEM_HOST) {
        // stage through the context's Merkle buffers: [leaves | nodes]
        lk.lock();
        const size_t need = (2 * n_leaves - 1) * 32;
        if (need > ctx->merkle_leaves_bytes) {
            if (ctx->merkle_leaves) cudaFree(ctx->merkle_leaves);
            ctx->merkle_leaves = nullptr;
            ctx->merkle_leaves_bytes = 0;
            if (cudaMalloc(&ctx->merkle_leaves, need) != cudaSuccess) {
                cudaGetLastError();
                return fail(PSD_ERR_NOMEM, "cudaMalloc(%zu) for the tree failed", need);
            }
            ctx->merkle_leaves_bytes = need;
        }
        uint32_t* base = static_cast<uint32_t*>(ctx->merkle_leaves);
        PSD_CUDA(cudaMemcpyAsync(base, leaves, n_leaves * 32, cudaMemcpyHostToDevice, s));
        d_leaves = base;
        d_nodes = base + n_leaves * 8;
    }
    // every level reads the previous one and writes its own slice of `nodes`, all in `repr`
    const uint32_t* src = d_leaves;
    uint32_t* dst = d_nodes;
    for (size_t pairs = n_leaves / 2; pairs >= 1; pairs /= 2) {
        SpongeArgs a = base_args(ctx, pairs);
        a.in = src;
        a.n_absorb = 2;
        a.in_layout = kLayoutAos;
        a.in_repr = repr;
        a.do_squeeze = 1;
        a.n_squeeze = 1;
        a.out = dst;
        a.out_layout = kLayoutAos;
        a.out_repr = repr;
        st = launch(ctx, a, s);
        if (st != PSD_OK) return st;
        src = dst;
        dst += pairs * 8;
        if (pairs == 1) break;
    }
    if (mem == PSD_MEM_DEVICE) return PSD_OK;   // asynchronous on the caller's stream
    PSD_CUDA(cudaMemcpyAsync(nodes, d_nodes, (n_leaves - 1) * 32, cudaMemcpyDeviceToHost, s));
    PSD_CUDA(cudaStreamSynchronize(s));
    return check_err_flag(ctx, s);
}

extern "C" psd_status psd_merkle_path(const uint64_t* leaves, const uint64_t* nodes, size_t n_leaves, size_t index,
                                      uint64_t* path) {
    if (!leaves || !nodes || !path) return fail(PSD_ERR_ARG, "null argument");
    if (n_leaves < 2 || (n_leaves & (n_leaves - 1)) || index >= n_leaves) return fail(PSD_ERR_ARG, "bad n_leaves/index");
    const uint64_t* level = leaves;   // current level, `width` elements
    size_t width = n_leaves, pos = index, next_off = 0;
    while (width > 1) {
        memcpy(path, level + (pos ^ 1) * 4, 32);
        path += 4;
        level = nodes + next_off * 4;
        next_off += width / 2;
        width /= 2;
        pos /= 2;
    }
    return PSD_OK;
}

// ------------------------------------------------------------------------------------------------
// Merkle root without any allocation (see include/poseidon_b200.h).  The level buffers live in a
// fixed array inside the library image instead of a cudaMalloc'd block: hosts that give all device
// memory to their own arena at start-up cannot tolerate a cudaMalloc on the request path.
// ------------------------------------------------------------------------------------------------
namespace {

constexpr size_t kNoAllocLevel1 = PSD_NOALLOC_MAX_LEAVES / 2;   // elements in the first level
constexpr size_t kNoAllocLevel2 = PSD_NOALLOC_MAX_LEAVES / 4;   // elements in the second level
// 8 words per element; 1.5 MiB on every GPU that loads the library, for the life of the process.
static __device__ __align__(256) uint32_t g_noalloc_scratch[(kNoAllocLevel1 + kNoAllocLevel2) * 8];

// In host code the name g_noalloc_scratch is only a handle for the symbol, not an address the GPU
// can use; the real device address is looked up once per device and cached here.
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_ptr[64] = {nullptr};

// Requires the target device to be current.
psd_status noalloc_scratch_for(int device, uint32_t** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(PSD_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d failed: %s", device,
                        cudaGetErrorString(e));
        }
        g_noalloc_ptr[device] = static_cast<uint32_t*>(p);
    }
    *out = g_noalloc_ptr[device];
    return PSD_OK;
}

}  // namespace

extern "C" psd_status psd_merkle_root_noalloc(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, uint64_t* root,
                                              int repr, void* stream) {
    if (!ctx) return fail(PSD_ERR_ARG, "null context");
    if (!leaves || !root) return fail(PSD_ERR_ARG, "null buffer");
    if (!valid_enum(repr, 1)) return fail(PSD_ERR_ARG, "bad repr value");
    if (ctx->rate < 2) return fail(PSD_ERR_UNSUPPORTED, "2-to-1 compression needs rate >= 2");
    if (n_leaves == 0 || (n_leaves & (n_leaves - 1))) return fail(PSD_ERR_ARG, "n_leaves must be a power of two");
    if (n_leaves > PSD_NOALLOC_MAX_LEAVES)
        return fail(PSD_ERR_ARG,
                    "psd_merkle_root_noalloc: n_leaves = %zu exceeds PSD_NOALLOC_MAX_LEAVES = %d (the fixed scratch "
                    "array holds no more); use psd_merkle_root or split the tree",
                    n_leaves, PSD_NOALLOC_MAX_LEAVES);
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return fail(PSD_ERR_CUDA, "cudaSetDevice(%d) failed", ctx->device);
    cudaStream_t s = (cudaStream_t)stream;
    if (n_leaves == 1) {
        PSD_CUDA(cudaMemcpyAsync(root, leaves, 32, cudaMemcpyDeviceToDevice, s));
        return PSD_OK;
    }
    uint32_t* scratch = nullptr;
    psd_status st = noalloc_scratch_for(ctx->device, &scratch);
    if (st != PSD_OK) return st;
    // Levels alternate between the two halves of the scratch array; the last level (one pair) is
    // written straight to the caller's root.  Every element a level reads was written by the level
    // before it in this same call, so stale contents from earlier calls are never observed.
    uint32_t* bufs[2] = {scratch, scratch + kNoAllocLevel1 * 8};
    const uint32_t* src = as_u32(leaves);
    int src_repr = repr;
    int which = 0;
    for (size_t pairs = n_leaves / 2; pairs >= 1; pairs /= 2) {
        SpongeArgs a = base_args(ctx, pairs);
        a.in = src;
        a.n_absorb = 2;
        a.in_layout = kLayoutAos;
        a.in_repr = src_repr;
        a.do_squeeze = 1;
        a.n_squeeze = 1;
        a.out = pairs == 1 ? as_u32(root) : bufs[which];
        a.out_layout = kLayoutAos;
        a.out_repr = pairs == 1 ? repr : kReprMontgomery;
        st = launch(ctx, a, s);
        if (st != PSD_OK) return st;
        src = bufs[which];
        src_repr = kReprMontgomery;
        which ^= 1;
        if (pairs == 1) break;
    }
    return PSD_OK;
}

// ------------------------------------------------------------------------------------------------
// Ragged batches: every sponge absorbs its own number of elements.
// ------------------------------------------------------------------------------------------------
extern "C" psd_status psd_absorb_squeeze_ragged_batch(psd_ctx* ctx, const uint64_t* inputs, const uint64_t* offsets, size_t n,
                                                      uint32_t n_squeeze, uint64_t* outputs, int repr, int mem, void* stream) {
    if (!ctx) return fail(PSD_ERR_ARG, "null context");
    if (!valid_enum(repr, 1) || !valid_enum(mem, 1)) return fail(PSD_ERR_ARG, "bad repr/mem value");
    if (n == 0) return PSD_OK;
    if (!offsets) return fail(PSD_ERR_ARG, "null offsets");
    if (n_squeeze && !outputs) return fail(PSD_ERR_ARG, "null output buffer");
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return fail(PSD_ERR_CUDA, "cudaSetDevice(%d) failed", ctx->device);
    auto fill = [&](SpongeArgs& a, const uint32_t* d_in, const uint64_t* d_off, uint64_t base, uint32_t* d_out) {
        a.in = d_in;
        a.in_offsets = d_off;
        a.in_offset_base = base;
        a.in_layout = kLayoutAos;
        a.in_repr = repr;
        a.do_squeeze = 1;
        a.n_squeeze = n_squeeze;
        a.out = d_out;
        a.out_layout = kLayoutAos;
        a.out_repr = repr;
    };
    if (mem == PSD_MEM_DEVICE) {
        // offsets and inputs are device pointers; nothing can be validated on the host
        if (!inputs) return fail(PSD_ERR_ARG, "null inputs");
        SpongeArgs a = base_args(ctx, n);
        fill(a, as_u32(inputs), offsets, 0, as_u32(outputs));
        return launch(ctx, a, (cudaStream_t)stream);
    }
    // host: validate the offsets, then stage sponge ranges of at most ~128 MiB of input each
    if (offsets[0] != 0) return fail(PSD_ERR_ARG, "offsets[0] must be 0");
    for (size_t i = 0; i < n; i++) {
        if (offsets[i + 1] < offsets[i]) return fail(PSD_ERR_ARG, "offsets must be non-decreasing (index %zu)", i + 1);
        if (offsets[i + 1] - offsets[i] > 0xffffffffull) return fail(PSD_ERR_ARG, "sponge %zu absorbs too many elements", i);
    }
    if (offsets[n] && !inputs) return fail(PSD_ERR_ARG, "null inputs");
    std::lock_guard<std::mutex> lk(ctx->stage_mu);
    const uint64_t max_el = (128ull << 20) / 32;
    const size_t max_sponges = 1u << 17;
    int slot = 0;
    psd_status st = PSD_OK;
    size_t first = 0;
    // chunk boundaries first, to size the staging buffers once
    std::vector<size_t> cuts{0};
    uint64_t need_in = 16, need_cnt = 1;
    while (first < n) {
        size_t last = first + 1;
        while (last < n && last - first < max_sponges && offsets[last + 1] - offsets[first] <= max_el) last++;
        need_in = std::max<uint64_t>(need_in, (offsets[last] - offsets[first]) * 32);
        need_cnt = std::max<uint64_t>(need_cnt, last - first);
        cuts.push_back(last);
        first = last;
    }
    // stage_in: [inputs][offsets (cnt + 1) u64]; stage_out: outputs
    const size_t in_bytes = (size_t)((need_in + 255) / 256 * 256 + (need_cnt + 1) * 8);
    st = ensure_staging(ctx, in_bytes, std::max<size_t>((size_t)need_cnt * n_squeeze * 32, 16));
    if (st != PSD_OK) return st;
    for (size_t c = 0; c + 1 < cuts.size(); c++, slot = (slot + 1) % kStageSlots) {
        const size_t lo = cuts[c], hi = cuts[c + 1], cnt = hi - lo;
        const uint64_t el = offsets[hi] - offsets[lo];
        cudaStream_t s = ctx->streams[slot];
        char* d_in = static_cast<char*>(ctx->stage_in[slot]);
        uint64_t* d_off = reinterpret_cast<uint64_t*>(d_in + (el * 32 + 255) / 256 * 256);
        if (el) PSD_CUDA(cudaMemcpyAsync(d_in, inputs + offsets[lo] * 4, el * 32, cudaMemcpyHostToDevice, s));
        PSD_CUDA(cudaMemcpyAsync(d_off, offsets + lo, (cnt + 1) * 8, cudaMemcpyHostToDevice, s));
        SpongeArgs a = base_args(ctx, cnt);
        fill(a, reinterpret_cast<const uint32_t*>(d_in), d_off, offsets[lo], static_cast<uint32_t*>(ctx->stage_out[slot]));
        st = launch(ctx, a, s);
        if (st != PSD_OK) return st;
        if (n_squeeze)
            PSD_CUDA(cudaMemcpyAsync(outputs + lo * n_squeeze * 4, ctx->stage_out[slot], cnt * (size_t)n_squeeze * 32,
                                     cudaMemcpyDeviceToHost, s));
    }
    for (int s = 0; s < kStageSlots; s++) PSD_CUDA(cudaStreamSynchronize(ctx->streams[s]));
    return check_err_flag(ctx, ctx->streams[0]);
}

// ------------------------------------------------------------------------------------------------
// One process, several GPUs.  Sponges never interact (a sponge is a plain value the caller spreads
// over workers, reference src/poseidon/mod.rs:44-60), so a batch shards into contiguous ranges, one per
// device, with no exchange at all; the Merkle tree has one exchange step: the subtree roots (32 bytes
// per device) are copied to the first device's own memory, where the top levels are hashed.
// ------------------------------------------------------------------------------------------------
#include <thread>

namespace {

constexpr int kMaxMultiGpus = 64;

// ctxs must be n_gpus distinct contexts of the same parameter set (same numbers; any device each).
psd_status check_multi(psd_ctx* const* ctxs, int n_gpus) {
    if (!ctxs) return fail(PSD_ERR_ARG, "null context array");
    if (n_gpus < 1 || n_gpus > kMaxMultiGpus) return fail(PSD_ERR_ARG, "n_gpus must be in [1, %d]", kMaxMultiGpus);
    for (int g = 0; g < n_gpus; g++) {
        if (!ctxs[g]) return fail(PSD_ERR_ARG, "null context at index %d", g);
        for (int h = 0; h < g; h++)
            if (ctxs[h] == ctxs[g]) return fail(PSD_ERR_ARG, "context %d and %d are the same object", h, g);
        const psd_ctx *a = ctxs[0], *b = ctxs[g];
        if (a->field != b->field || a->t != b->t || a->rate != b->rate || a->full_rounds != b->full_rounds ||
            a->partial_rounds != b->partial_rounds || a->alpha != b->alpha ||
            memcmp(a->ark_mont.data(), b->ark_mont.data(), a->ark_mont.size() * sizeof(Fe)) != 0 ||
            memcmp(a->mds_mont.data(), b->mds_mont.data(), a->mds_mont.size() * sizeof(Fe)) != 0)
            return fail(PSD_ERR_ARG, "context %d holds a different parameter set than context 0", g);
    }
    return PSD_OK;
}

// Runs work(g) for every device on its own host thread (each drives its own device's streams) and
// returns the first failure with its message.
template <class Work>
psd_status run_per_device(int n_gpus, Work work) {
    std::vector<psd_status> st((size_t)n_gpus, PSD_OK);
    std::vector<std::string> msg((size_t)n_gpus);
    auto body = [&](int g) {
        st[(size_t)g] = work(g);
        if (st[(size_t)g] != PSD_OK) msg[(size_t)g] = t_last_error;   // thread-local: copy it out
    };
    std::vector<std::thread> threads;
    for (int g = 1; g < n_gpus; g++) threads.emplace_back(body, g);
    body(0);
    for (auto& th : threads) th.join();
    for (int g = 0; g < n_gpus; g++)
        if (st[(size_t)g] != PSD_OK) return fail(st[(size_t)g], "device %d of %d: %s", g, n_gpus, msg[(size_t)g].c_str());
    return PSD_OK;
}

// Contiguous shard of rank g: the first n % world ranks get one extra unit.
void shard_range(size_t n, int g, int world, size_t* lo, size_t* hi) {
    const size_t base = n / (size_t)world, extra = n % (size_t)world;
    *lo = (size_t)g * base + std::min<size_t>((size_t)g, extra);
    *hi = *lo + base + ((size_t)g < extra ? 1 : 0);
}

psd_status staged_shard(psd_ctx* ctx, const uint64_t* inputs, size_t n, uint32_t n_absorb, uint32_t n_squeeze,
                        uint64_t* outputs, bool permute, int layout, int repr, size_t lo, size_t hi) {
    if (hi <= lo) return PSD_OK;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return fail(PSD_ERR_CUDA, "cudaSetDevice(%d) failed", ctx->device);
    if (permute) {
        auto fill = [&](SpongeArgs& a, void* d_in, void*) {
            a.state_in = (const uint32_t*)d_in;
            a.state_in_layout = layout;
            a.state_in_repr = repr;
            a.state_out = (uint32_t*)d_in;
            a.state_out_layout = layout;
            a.state_out_repr = repr;
            a.n_raw = 1;
        };
        StagedCall c{n, inputs, ctx->t, outputs, ctx->t, 0, layout, true};
        c.first0 = lo, c.count = hi - lo;
        return run_staged(ctx, c, fill);
    }
    auto fill = [&](SpongeArgs& a, void* d_in, void* d_out) {
        a.in = (const uint32_t*)d_in;
        a.n_absorb = n_absorb;
        a.in_layout = layout;
        a.in_repr = repr;
        a.do_squeeze = 1;
        a.n_squeeze = n_squeeze;
        a.out = (uint32_t*)d_out;
        a.out_layout = layout;
        a.out_repr = repr;
    };
    StagedCall c{n, inputs, n_absorb, outputs, n_squeeze, 0, layout, false};
    c.first0 = lo, c.count = hi - lo;
    return run_staged(ctx, c, fill);
}

}  // namespace

extern "C" {

void psd_shard_range(size_t n, int rank, int world, size_t* lo, size_t* hi) {
    size_t a = 0, b = 0;
    if (world >= 1 && rank >= 0 && rank < world) shard_range(n, rank, world, &a, &b);
    if (lo) *lo = a;
    if (hi) *hi = b;
}

psd_status psd_permute_batch_multi(psd_ctx* const* ctxs, int n_gpus, uint64_t* states, size_t n, int layout, int repr) {
    psd_status st = check_multi(ctxs, n_gpus);
    if (st != PSD_OK) return st;
    st = check_batch_args(ctxs[0], n ? (void*)states : (void*)ctxs, layout, repr, PSD_MEM_HOST);
    if (st != PSD_OK || n == 0) return st;
    return run_per_device(n_gpus, [&](int g) {
        size_t lo, hi;
        shard_range(n, g, n_gpus, &lo, &hi);
        return staged_shard(ctxs[g], states, n, 0, 0, states, true, layout, repr, lo, hi);
    });
}

psd_status psd_absorb_squeeze_batch_multi(psd_ctx* const* ctxs, int n_gpus, const uint64_t* inputs, size_t n,
                                          uint32_t n_absorb, uint32_t n_squeeze, uint64_t* outputs, int layout, int repr) {
    psd_status st = check_multi(ctxs, n_gpus);
    if (st != PSD_OK) return st;
    st = check_batch_args(ctxs[0], (n && n_absorb) ? (const void*)inputs : (const void*)ctxs, layout, repr, PSD_MEM_HOST);
    if (st != PSD_OK) return st;
    if (n && n_squeeze && !outputs) return fail(PSD_ERR_ARG, "null output buffer");
    if (n == 0) return PSD_OK;
    return run_per_device(n_gpus, [&](int g) {
        size_t lo, hi;
        shard_range(n, g, n_gpus, &lo, &hi);
        return staged_shard(ctxs[g], inputs, n, n_absorb, n_squeeze, outputs, false, layout, repr, lo, hi);
    });
}

psd_status psd_merkle_root_multi(psd_ctx* const* ctxs, int n_gpus, const uint64_t* const* leaves_per_gpu,
                                 size_t leaves_per_gpu_n, uint64_t* root, int repr, int mem, void* const* streams) {
    psd_status st = check_multi(ctxs, n_gpus);
    if (st != PSD_OK) return st;
    if (!leaves_per_gpu || !root) return fail(PSD_ERR_ARG, "null argument");
    if (!valid_enum(repr, 1) || !valid_enum(mem, 1)) return fail(PSD_ERR_ARG, "bad repr/mem value");
    for (int g = 0; g < n_gpus; g++)
        if (!leaves_per_gpu[g]) return fail(PSD_ERR_ARG, "null leaf pointer for device %d", g);
    if (ctxs[0]->rate < 2) return fail(PSD_ERR_UNSUPPORTED, "2-to-1 compression needs rate >= 2");
    const size_t L = leaves_per_gpu_n;
    if (L == 0 || (L & (L - 1))) return fail(PSD_ERR_ARG, "leaves per device must be a power of two");
    if (n_gpus & (n_gpus - 1)) return fail(PSD_ERR_ARG, "n_gpus must be a power of two");
    if (n_gpus == 1) return psd_merkle_root(ctxs[0], leaves_per_gpu[0], L, root, repr, mem, streams ? streams[0] : nullptr);

    psd_ctx* c0 = ctxs[0];
    // every context's Merkle buffers are held for the whole call
    std::vector<std::unique_lock<std::mutex>> locks;
    for (int g = 0; g < n_gpus; g++) locks.emplace_back(ctxs[g]->merkle_mu);
    // device 0: room for the gathered subtree roots + the final root, and one event per device
    {
        DeviceGuard guard(c0->device);
        if (!guard.ok) return fail(PSD_ERR_CUDA, "cudaSetDevice(%d) failed", c0->device);
        if (!c0->multi_roots) PSD_CUDA(cudaMalloc(&c0->multi_roots, (size_t)(kMaxMultiGpus + 1) * 32));
    }
    uint32_t* gathered = static_cast<uint32_t*>(c0->multi_roots);          // [n_gpus] elements, device 0
    uint32_t* d_final = gathered + (size_t)kMaxMultiGpus * 8;               // the root, device 0
    auto stream_of = [&](int g) {
        return (mem == PSD_MEM_DEVICE && streams) ? (cudaStream_t)streams[g] : ctxs[g]->streams[0];
    };
    // Stage 1, every device on its own thread and stream: [upload] -> subtree levels -> the subtree root
    // is copied into device 0's buffer (a plain peer copy; no kernel reads another GPU's memory) -> event.
    // With one leaf per device there is no subtree: the leaves themselves are gathered, in `repr`.
    const int gathered_repr = L == 1 ? repr : kReprMontgomery;
    st = run_per_device(n_gpus, [&](int g) -> psd_status {
        psd_ctx* ctx = ctxs[g];
        DeviceGuard guard(ctx->device);
        if (!guard.ok) return fail(PSD_ERR_CUDA, "cudaSetDevice(%d) failed", ctx->device);
        cudaStream_t s = stream_of(g);
        if (!ctx->multi_event) PSD_CUDA(cudaEventCreateWithFlags(&ctx->multi_event, cudaEventDisableTiming));
        if (ctx->device != c0->device) {
            // direct DMA for the 32-byte copy where the devices can reach each other (NVLink); without it the
            // runtime stages the copy through the host, which is only slower.  Once per device pair.
            static std::atomic<uint64_t> tried[64];
            const uint64_t bit = 1ull << c0->device;
            if (!(tried[ctx->device].fetch_or(bit) & bit)) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, ctx->device, c0->device) == cudaSuccess && can)
                    cudaDeviceEnablePeerAccess(c0->device, 0);
                cudaGetLastError();   // "already enabled" is fine
            }
        }
        const uint32_t* src = as_u32(leaves_per_gpu[g]);
        if (mem == PSD_MEM_HOST) {
            psd_status r = merkle_upload_leaves(ctx, leaves_per_gpu[g], L, s, &src);
            if (r != PSD_OK) return r;
        }
        const uint32_t* sub = src;
        if (L > 1) {
            psd_status r = merkle_reduce_async(ctx, src, repr, L, nullptr, kReprMontgomery, s, &sub);
            if (r != PSD_OK) return r;
        }
        PSD_CUDA(cudaMemcpyPeerAsync(gathered + (size_t)g * 8, c0->device, sub, ctx->device, 32, s));
        PSD_CUDA(cudaEventRecord(ctx->multi_event, s));
        return PSD_OK;
    });
    // Stage 2, device 0: wait for every gather, hash the n_gpus local roots, deliver.
    psd_status result = st;
    DeviceGuard guard(c0->device);
    cudaStream_t s0 = stream_of(0);
    do {
        if (result != PSD_OK) break;
        for (int g = 1; g < n_gpus; g++) {
            if (cudaStreamWaitEvent(s0, ctxs[g]->multi_event, 0) != cudaSuccess) {
                result = fail(PSD_ERR_CUDA, "cudaStreamWaitEvent failed: %s", cudaGetErrorString(cudaGetLastError()));
                break;
            }
        }
        if (result != PSD_OK) break;
        // the top tree uses the upper part of device 0's scratch as a fresh context would: its own levels fit
        // in (n_gpus/2 + n_gpus/4) elements, far below what the subtree stage already allocated; but the
        // subtree stage of device 0 is complete in stream order before this runs, so reuse is safe
        uint32_t* dst = mem == PSD_MEM_DEVICE ? as_u32(root) : d_final;
        result = merkle_reduce_async(c0, gathered, gathered_repr, (size_t)n_gpus, dst, repr, s0, nullptr);
        if (result != PSD_OK) break;
        if (mem == PSD_MEM_HOST && cudaMemcpyAsync(root, d_final, 32, cudaMemcpyDeviceToHost, s0) != cudaSuccess) {
            result = fail(PSD_ERR_CUDA, "root download failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        if (cudaStreamSynchronize(s0) != cudaSuccess) {
            result = fail(PSD_ERR_CUDA, "top levels failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
    } while (0);
    // all streams idle before the buffers are released; collect the non-canonical flags of every device
    for (int g = 0; g < n_gpus; g++) {
        DeviceGuard dg(ctxs[g]->device);
        cudaStreamSynchronize(stream_of(g));
        if (result == PSD_OK && repr == PSD_REPR_CANONICAL) result = check_err_flag(ctxs[g], stream_of(g));
    }
    return result;
}

}  // extern "C"
