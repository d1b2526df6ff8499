/* poseidon_b200.h -- C ABI of the B200-native batched Poseidon sponge engine.
 *
 * This is the drop-in boundary for ONE path of arkworks-rs/sponge: the Poseidon permutation and
 * the duplex absorb/squeeze bookkeeping of `PoseidonSponge<F>` (reference: src/poseidon/mod.rs),
 * lifted to N independent sponges that advance in lock-step on one GPU.  A Rust FFI crate
 * (`impl CryptographicSponge for B200PoseidonSponge<F>`), the C++ wrapper
 * include/poseidon_b200.hpp and the Python ctypes binding sponge_b200/ all bind exactly these
 * symbols.  There is no CPU fallback behind any of them: every call that computes runs the
 * sm_100a kernels or returns PSD_ERR_CUDA.
 *
 * Element format.  One field element is 32 bytes = 4 little-endian u64 limbs, which is the memory
 * image of `ark_ff::Fp<MontBackend<_,4>,4>` (a Rust `&[Fr]` is an AoS array of these).
 *   PSD_REPR_MONTGOMERY : the limbs hold x*2^256 mod p  (what `Fr.0.0` holds; zero-cost for Rust)
 *   PSD_REPR_CANONICAL  : the limbs hold x              (`into_bigint()`); inputs must be < p
 * Layout of a batch of n sponges with w elements each:
 *   PSD_LAYOUT_AOS : elem(i,k) at u64 offset (i*w + k)*4                       (Rust slice order)
 *   PSD_LAYOUT_SOA : elem(i,k) is split in two 16-byte halves h=0,1 stored at
 *                    u64 offset ((k*2 + h)*n + i)*2  -- a warp reads 32 consecutive 16-byte words
 * State lanes follow the reference: capacity lanes first, rate lanes at [capacity + j]
 * (mod.rs:128,143,159).
 *
 * Memory spaces.  PSD_MEM_DEVICE: pointers are device pointers on the context's device and the
 * call is asynchronous on `stream`.  PSD_MEM_HOST: pointers are host memory (pinned or pageable);
 * the library stages through device buffers it owns, overlapping H2D, kernel and D2H, and returns
 * when the outputs are in host memory.
 *
 * Threading: a psd_ctx is immutable after creation and may be shared between threads; batch calls
 * with PSD_MEM_HOST serialise on the context's staging buffers.  psd_batch / psd_sponge handles
 * are not thread-safe (same as `&mut self` in the reference).
 *
 * Errors: every function returns a psd_status; nothing aborts or throws across the ABI
 * (the reference panics: shape asserts mod.rs:196-203, unwrap absorb.rs:163).
 */
#ifndef POSEIDON_B200_H
#define POSEIDON_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum psd_status {
    PSD_OK = 0,
    PSD_ERR_ARG = 1,          /* null pointer, unknown enum value, n too large */
    PSD_ERR_SHAPE = 2,        /* PoseidonConfig::new shape asserts (mod.rs:196-203) would fire */
    PSD_ERR_NONCANONICAL = 3, /* a canonical input is >= p (ark-ff from_bigint -> None) */
    PSD_ERR_CUDA = 4,         /* CUDA runtime error; see psd_last_error_string */
    PSD_ERR_UNSUPPORTED = 5,  /* e.g. state width outside [2,9], rate < 2 for Merkle */
    PSD_ERR_NOMEM = 6         /* host/device allocation failed */
} psd_status;

typedef enum psd_field { PSD_FIELD_BN254_FR = 0, PSD_FIELD_BLS12_381_FR = 1 } psd_field;
typedef enum psd_repr { PSD_REPR_CANONICAL = 0, PSD_REPR_MONTGOMERY = 1 } psd_repr;
typedef enum psd_layout { PSD_LAYOUT_AOS = 0, PSD_LAYOUT_SOA = 1 } psd_layout;
typedef enum psd_mem { PSD_MEM_HOST = 0, PSD_MEM_DEVICE = 1 } psd_mem;
/* DuplexSpongeMode, reference src/lib.rs:198-210 */
typedef enum psd_mode { PSD_MODE_ABSORBING = 0, PSD_MODE_SQUEEZING = 1 } psd_mode;

/* Mirrors `PoseidonConfig<F>` field for field (reference src/poseidon/mod.rs:23-42; the 0.3.0
 * release called it PoseidonParameters).  ark is indexed [round][lane], mds [row][col]; each entry
 * is 4 u64 limbs in `repr`. */
typedef struct psd_params {
    int32_t field;            /* psd_field */
    uint32_t full_rounds;
    uint32_t partial_rounds;
    uint64_t alpha;
    uint32_t rate;
    uint32_t capacity;
    const uint64_t* ark;      /* (full_rounds+partial_rounds) * (rate+capacity) * 4 u64 */
    const uint64_t* mds;      /* (rate+capacity)^2 * 4 u64 */
    int32_t repr;             /* psd_repr of ark/mds */
} psd_params;

typedef struct psd_ctx psd_ctx;       /* parameters resident on one GPU */
typedef struct psd_batch psd_batch;   /* n sponges in lock-step, states resident in HBM */
typedef struct psd_sponge psd_sponge; /* one PoseidonSponge<F> (a batch of 1) */

/* ---- library ---------------------------------------------------------------------------------- */
const char* psd_version(void);
/* Thread-local text of the last error raised on the calling thread. */
const char* psd_last_error_string(void);
const char* psd_status_string(psd_status s);
/* Number of visible CUDA devices (0 if the runtime cannot be initialised). */
int psd_device_count(void);

/* ---- parameters -------------------------------------------------------------------------------
 * psd_ctx_create replaces PoseidonConfig::new + the per-sponge `parameters.clone()` of
 * PoseidonSponge::new (mod.rs:187-213, 226): shapes are validated once, constants are converted to
 * Montgomery form and uploaded to the device's constant bank once. */
psd_status psd_ctx_create(const psd_params* params, int device, psd_ctx** out);
void psd_ctx_destroy(psd_ctx* ctx);
/* Read back scalar fields of the context's parameter set (ark/mds pointers are NULL). */
psd_status psd_ctx_get_params(const psd_ctx* ctx, psd_params* out);

/* Host-only parameter generation, replaces find_poseidon_ark_and_mds (traits.rs:105-146) with the
 * Grain LFSR of grain_lfsr.rs:15-189.  Writes canonical values: ark (full+partial)*(rate+1)*4 u64,
 * mds (rate+1)^2*4 u64.  prime_bits must equal the field's modulus size (254 / 255). */
psd_status psd_find_poseidon_ark_and_mds(int32_t field, uint64_t prime_bits, uint32_t rate,
                                         uint32_t full_rounds, uint32_t partial_rounds,
                                         uint32_t skip_matrices, uint64_t* ark, uint64_t* mds);
/* Replaces PoseidonDefaultConfigField::get_default_poseidon_parameters (traits.rs:59-102) for the
 * default tables of the reference's BLS12-381 Fr (test.rs:13-32); for BN254 Fr the same table
 * shape is offered with the round numbers of BASELINE.json (rate 2: 8/57, rate 4: 8/60, alpha 5).
 * Fills the scalar fields of *out and writes ark/mds (canonical) into caller buffers sized for
 * the worst case: ark 73*9*4 u64, mds 81*4 u64.  Returns PSD_ERR_UNSUPPORTED where the reference
 * returns None. */
psd_status psd_default_params(int32_t field, uint32_t rate, int optimized_for_weights,
                              psd_params* out, uint64_t* ark_buf, uint64_t* mds_buf);

/* ---- representation / layout helpers (ark-ff into_bigint / from_bigint; absorb.rs:106-120) ----
 * In place over n elements, on the context's device. */
psd_status psd_to_montgomery(psd_ctx* ctx, uint64_t* elems, size_t n, int mem, void* stream);
psd_status psd_from_montgomery(psd_ctx* ctx, uint64_t* elems, size_t n, int mem, void* stream);

/* ---- batch entry points (new) --------------------------------------------------
 * psd_permute_batch: n independent applications of PoseidonSponge::permute (mod.rs:95-118), in
 * place; each state is rate+capacity elements. */
psd_status psd_permute_batch(psd_ctx* ctx, uint64_t* states, size_t n, int layout, int repr, int mem,
                             void* stream);

/* For each of n sponges: PoseidonSponge::new; absorb(&inputs[i][0..n_absorb]);
 * squeeze_native_field_elements(n_squeeze)  (mod.rs:219-254, 320-342).  inputs: n*n_absorb
 * elements, outputs: n*n_squeeze elements, both in `layout`/`repr`. */
psd_status psd_absorb_squeeze_batch(psd_ctx* ctx, const uint64_t* inputs, size_t n, uint32_t n_absorb,
                                    uint32_t n_squeeze, uint64_t* outputs, int layout, int repr,
                                    int mem, void* stream);

/* Same, ending in squeeze_bytes(n_bytes) (mod.rs:256-270): outputs is n*n_bytes bytes, sponge i at
 * byte offset i*n_bytes. */
psd_status psd_absorb_squeeze_bytes_batch(psd_ctx* ctx, const uint64_t* inputs, size_t n,
                                          uint32_t n_absorb, uint32_t n_bytes, uint8_t* outputs,
                                          int layout, int repr, int mem, void* stream);

/* Ragged form: sponge i absorbs the elements inputs[offsets[i] .. offsets[i+1]) (AoS, concatenated;
 * offsets has n+1 entries, offsets[0] = 0, non-decreasing; a sponge may absorb nothing) and then squeezes
 * n_squeeze elements -- what the reference's callers do when they absorb lists of differing lengths into
 * different sponges (src/poseidon/tests.rs:45-69; AbsorbWithLength, absorb.rs:82-101).  Same kernel variant
 * as the lock-step calls, in a form that tolerates differing lengths; sort or bucket by length on the
 * caller's side if the lengths vary widely (with the default kernel a thread block of up to 640 sponges, with
 * the integer kernels a warp, takes as long as its longest sponge).  outputs: n*n_squeeze
 * elements.  mem = PSD_MEM_DEVICE: inputs, offsets and outputs are device pointers. */
psd_status psd_absorb_squeeze_ragged_batch(psd_ctx* ctx, const uint64_t* inputs, const uint64_t* offsets, size_t n,
                                           uint32_t n_squeeze, uint64_t* outputs, int repr, int mem, void* stream);

/* Merkle root over n_leaves (a power of two) with the 2-to-1 compression
 *   H(l, r) := { s = PoseidonSponge::new; s.absorb(&[l, r]); s.squeeze_native_field_elements(1)[0] }
 * (rate >= 2).  leaves: n_leaves elements (AoS); root: 1 element.  Intermediate levels live in a
 * device scratch buffer (n_leaves/2 + n_leaves/4 elements) that the call allocates with cudaMalloc
 * when it is missing or too small and the context keeps until it is destroyed; calls on one
 * context are serialised.  The call returns when the root has been written. */
psd_status psd_merkle_root(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, uint64_t* root,
                           int repr, int mem, void* stream);
/* The whole tree: like psd_merkle_root but every internal node is returned.  nodes receives
 * n_leaves - 1 elements, bottom-up: the n_leaves/2 parents of the leaves first, then the n_leaves/4
 * nodes above them, ..., the root last (node j of level k, k = 1 for the leaves' parents, is at
 * element offset n_leaves - n_leaves/2^(k-1) + j).  n_leaves >= 2, a power of two. */
psd_status psd_merkle_tree(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, uint64_t* nodes, int repr, int mem,
                           void* stream);
/* Authentication path of leaf `index` out of the arrays above (host pointers, pure indexing):
 * path receives log2(n_leaves) elements, the sibling at every level from the leaf's own up. */
psd_status psd_merkle_path(const uint64_t* leaves, const uint64_t* nodes, size_t n_leaves, size_t index, uint64_t* path);
/* One tree level only: out[i] = H(in[2i], in[2i+1]), i < n_pairs (device or host pointers). */
psd_status psd_compress_pairs(psd_ctx* ctx, const uint64_t* in, size_t n_pairs, uint64_t* out, int repr,
                              int mem, void* stream);

/* Merkle root for callers that must not allocate.  No counterpart in the reference (like
 * psd_merkle_root it is defined through the reference sponge; this variant exists only because of
 * how some hosts manage GPU memory).
 *
 * Who needs it: a proving service that carves up its whole device-memory budget at start-up and
 * hands every byte to its own arena cannot let a library call cudaMalloc afterwards -- the call can
 * fail outright (nothing is left outside the arena), it takes a process-wide driver lock that stalls
 * the other streams, and in such code bases an allocation on the request path is treated as a bug.
 * psd_merkle_root allocates its level buffers with cudaMalloc whenever a tree is larger than any
 * seen before on that context, so it is unusable there.
 *
 * This entry point computes the same root with the intermediate levels kept in a fixed scratch
 * array that is part of the library image: PSD_NOALLOC_MAX_LEAVES/2 + PSD_NOALLOC_MAX_LEAVES/4
 * elements = 1.5 MiB, resident on every GPU that loads the library until the process exits.  It
 * calls no cudaMalloc/cudaFree and no allocating library routine.  `leaves` (n_leaves elements,
 * AoS) and `root` (1 element) are DEVICE pointers on the context's device; the work is queued on
 * `stream` and the call returns without waiting (a non-canonical leaf is reported by the next
 * psd_stream_synchronize, as for the other device-pointer calls).
 *   - n_leaves must be a power of two and <= PSD_NOALLOC_MAX_LEAVES, else PSD_ERR_ARG.
 *   - The first call on a device looks up the scratch array's device address (once per device, with
 *     that device current); make that first call before any stream capture begins.
 *   - All calls share the one array and it keeps whatever the last call left in it.  Nothing has
 *     to be cleared: each level is written in full before the next level reads it.  Two calls that
 *     overlap on the GPU (different streams, or threads) race on it -- serialise them. */
#define PSD_NOALLOC_MAX_LEAVES 65536
psd_status psd_merkle_root_noalloc(psd_ctx* ctx, const uint64_t* leaves, size_t n_leaves, uint64_t* root, int repr,
                                   void* stream);

/* ---- one process, several GPUs ------------------------------------------------------------------
 * No counterpart in the reference, which leaves spreading independent sponges over workers to the
 * caller (a sponge is a plain value: poseidon/mod.rs:44-60).  ctxs[0..n_gpus) are contexts created from
 * the SAME psd_params on different devices (checked).  The batch calls take HOST buffers holding the
 * whole batch and give device g the contiguous range psd_shard_range(n, g, n_gpus): every device runs
 * its own upload / kernel / download pipeline from its own host thread, there is no exchange between
 * devices, and the calls return when all outputs are in host memory.  Layout/repr as in the
 * single-device calls (SoA shards are strided sub-blocks of the caller's arrays). */
void psd_shard_range(size_t n, int rank, int world, size_t* lo, size_t* hi);
psd_status psd_permute_batch_multi(psd_ctx* const* ctxs, int n_gpus, uint64_t* states, size_t n, int layout,
                                   int repr);
psd_status psd_absorb_squeeze_batch_multi(psd_ctx* const* ctxs, int n_gpus, const uint64_t* inputs, size_t n,
                                          uint32_t n_absorb, uint32_t n_squeeze, uint64_t* outputs, int layout,
                                          int repr);
/* Merkle root of the tree whose leaves are leaves_per_gpu[0] || leaves_per_gpu[1] || ... (n_gpus and
 * leaves_per_gpu_n powers of two; same H as psd_merkle_root).  Device g reduces its own subtree on its
 * own stream; the only exchange is the n_gpus subtree roots (32 bytes each), copied with
 * cudaMemcpyPeerAsync into a buffer in the first device's OWN memory, where that device hashes the top
 * log2(n_gpus) levels.  mem = PSD_MEM_DEVICE: leaves_per_gpu[g] is a device pointer on ctxs[g]'s device,
 * root a device pointer on ctxs[0]'s device, and streams (NULL, or n_gpus CUDA streams, one per device)
 * are the streams the leaves were produced on.  mem = PSD_MEM_HOST: all host pointers (they may point
 * into one contiguous array).  Returns when the root has been written. */
psd_status psd_merkle_root_multi(psd_ctx* const* ctxs, int n_gpus, const uint64_t* const* leaves_per_gpu,
                                 size_t leaves_per_gpu_n, uint64_t* root, int repr, int mem,
                                 void* const* streams);

/* ---- lock-step batch of sponges (absorb/squeeze bookkeeping lifted to batch) --------------------
 * A psd_batch is n PoseidonSponge<F> values that always receive calls of identical shape, so they
 * share one DuplexSpongeMode; states stay in HBM between calls (SoA, Montgomery). */
psd_status psd_batch_new(psd_ctx* ctx, size_t n, psd_batch** out);      /* PoseidonSponge::new x n */
psd_status psd_batch_clone(const psd_batch* b, psd_batch** out);        /* Clone (mod.rs:44) */
void psd_batch_free(psd_batch* b);
/* absorb(&[F; k]) on every sponge (mod.rs:232-254). elems: n*k elements. k == 0 is a no-op. */
psd_status psd_batch_absorb(psd_batch* b, const uint64_t* elems, uint32_t k, int layout, int repr,
                            int mem);
/* squeeze_native_field_elements(k) on every sponge (mod.rs:320-342). out: n*k elements. */
psd_status psd_batch_squeeze_native(psd_batch* b, uint32_t k, uint64_t* out, int layout, int repr,
                                    int mem);
/* squeeze_bytes(n_bytes) on every sponge (mod.rs:256-270). out: n*n_bytes bytes. */
psd_status psd_batch_squeeze_bytes(psd_batch* b, uint32_t n_bytes, uint8_t* out, int mem);
/* SpongeExt::into_state / from_state (lib.rs:188-195, mod.rs:344-367): states are n*(rate+capacity)
 * elements; mode/index are shared by the batch. */
psd_status psd_batch_get_state(const psd_batch* b, uint64_t* states, int layout, int repr, int mem,
                               int* mode, uint32_t* index);
psd_status psd_batch_set_state(psd_batch* b, const uint64_t* states, int layout, int repr, int mem,
                               int mode, uint32_t index);
size_t psd_batch_size(const psd_batch* b);
/* Run this batch's kernels and copies on the caller's stream (NULL = back to the batch's own
 * stream).  Calls still return only when their work is complete. */
psd_status psd_batch_set_stream(psd_batch* b, void* stream);

/* ---- single-sponge drop-in (CryptographicSponge, lib.rs:105-158; FieldBasedCryptographicSponge,
 * lib.rs:162-183).  A batch of one: latency-bound, exists for API parity. Host pointers. ---------- */
psd_status psd_sponge_new(psd_ctx* ctx, psd_sponge** out);                      /* new, mod.rs:219-230 */
psd_status psd_sponge_clone(const psd_sponge* s, psd_sponge** out);             /* Clone */
void psd_sponge_free(psd_sponge* s);
psd_status psd_sponge_absorb(psd_sponge* s, const uint64_t* elems, size_t n, int repr);   /* mod.rs:232-254 */
/* absorb(&[u8]) -- u64-LE length prefix, 31-byte LE chunks (absorb.rs:122-140) */
psd_status psd_sponge_absorb_bytes(psd_sponge* s, const uint8_t* bytes, size_t n);
psd_status psd_sponge_squeeze_native(psd_sponge* s, size_t n, uint64_t* out, int repr);   /* mod.rs:320-342 */
psd_status psd_sponge_squeeze_bytes(psd_sponge* s, size_t n_bytes, uint8_t* out);         /* mod.rs:256-270 */
/* squeeze_bits (mod.rs:272-286): one byte (0/1) per bit */
psd_status psd_sponge_squeeze_bits(psd_sponge* s, size_t n_bits, uint8_t* out);
/* squeeze_field_elements::<F2> for a non-native target field (mod.rs:288-317, lib.rs:61-100):
 * out = n elements of `target_field`, canonical. Same field -> native squeeze. */
psd_status psd_sponge_squeeze_field_elements(psd_sponge* s, int32_t target_field, size_t n, uint64_t* out);
/* fork(domain), lib.rs:149-157 */
psd_status psd_sponge_fork(const psd_sponge* s, const uint8_t* domain, size_t n, psd_sponge** out);
/* SpongeExt, lib.rs:188-195 */
psd_status psd_sponge_get_state(const psd_sponge* s, uint64_t* state, int repr, int* mode, uint32_t* index);
psd_status psd_sponge_set_state(psd_sponge* s, const uint64_t* state, int repr, int mode, uint32_t index);

/* ---- host/device memory helpers for callers without a CUDA runtime binding --------------------- */
psd_status psd_host_alloc(size_t bytes, void** out);   /* pinned host memory */
void psd_host_free(void* p);
psd_status psd_device_alloc(psd_ctx* ctx, size_t bytes, void** out);
void psd_device_free(psd_ctx* ctx, void* p);
psd_status psd_memcpy_h2d(psd_ctx* ctx, void* dst, const void* src, size_t bytes, void* stream);
psd_status psd_memcpy_d2h(psd_ctx* ctx, void* dst, const void* src, size_t bytes, void* stream);
psd_status psd_stream_synchronize(psd_ctx* ctx, void* stream);

/* Synthetic workload generator: element k of stream `seed` = four splitmix64 words
 * -> 256-bit LE, masked to the modulus size, redrawn until < p; written in `repr` (AoS) to a
 * device buffer.  Used by bench.py so inputs are identical on host and device. */
psd_status psd_synth_fill(psd_ctx* ctx, uint64_t* dev_elems, size_t n, uint64_t seed, uint64_t first_index,
                          int repr, void* stream);

/* ---- measurement hooks -------------------------------------------------------------------------- */
/* Number of kernels this library has launched in the calling process (monotonic). */
uint64_t psd_kernel_launch_count(void);
/* K0: sustained integer-multiply throughput of the device the context lives on.  Runs a
 * IMAD.WIDE.U32 (kind 0), IMAD lo (kind 1), the library's own Montgomery product (kind 2) or
 * squaring (kind 3) loop on every SM and returns results per second. */
psd_status psd_imad_peak(psd_ctx* ctx, int kind, double* per_second, double* elapsed_ms);
/* Select the permutation kernel variant for subsequent calls on this context (same results):
 * 0 = dense (ark -> S-box -> dense MDS every round, the reference's operation order),
 * 1 = optimised, fully unrolled rounds,
 * 2 = optimised, compact rolled phases,
 * 3 = 2 + products by round constants through digit tables (t = 3; 4 = the t = 5 form: lane updates only),
 * 5 = integer S-box + the whole linear layer on the FP64 pipe in 52-bit limbs (BN254 Fr, t = 2, 3, 5,
 *     alpha 5 or 17),
 * 6 = integer S-box + the whole linear layer on the tensor cores: every product by a round constant is a
 *     u8 x u8 -> s32 matrix product against the constant's 32 x 32 byte-digit table (mma.sync.m16n8k32),
 *     followed by one carry sweep and one 32-bit Montgomery row per output value (t = 2..5, alpha 5 or
 *     17, both fields).  Lock-step calls only: ragged batches run the next variant down.
 * 7 = the same contraction through tcgen05.mma.kind::i8 with the accumulators in tensor memory, and with
 *     every Montgomery reduction of the S-box done the same way (a reduction is a product by constants):
 *     the integer pipe only computes unreduced squares and products.  t = 2..5, alpha 5 or 17, both
 *     fields; the default.  Ragged batches too (the block permutes while any of its sponges
 *     needs to).  Where it has no kernel the request degrades to 6 (BN254)
 *     or to the integer variants (BLS12-381, where 6 measured slower than 3).
 * A request is served by the best kernel <= it that exists for the context's field, width and home of
 * constants; psd_ctx_active_variant / psd_ctx_describe report what a launch really runs.
 *
 * Where the constants live.  The kernels read round constants either from the 64 KiB constant bank
 * (one pool per (field, width) plus one for variant 5; fastest) or, in their shared-memory form, from a
 * copy that every thread block makes of the context's image in global memory.  psd_ctx_create looks for
 * an identical live parameter set on the device first (such contexts share one image), then for room in
 * the bank, then falls back to the shared-memory form, which any number of parameter sets can use at
 * once -- so creating a context never fails for lack of constant-bank space and a second parameter set
 * of the same shape runs the same algorithm as the first (measured within a few percent; the t = 5
 * variant-5 tables, 100 KiB, exist in the shared-memory form only).  Variants 6 and 7 keep their tables (1 or
 * 2 KiB per constant, 250 / 560 KiB for t = 3) in global memory, where the L1/L2 caches serve them (variant 7
 * stages a pair of rounds' tables into shared memory one pair ahead), and stage only the round-0
 * constants; any number of parameter sets can use them at once. */
psd_status psd_ctx_set_variant(psd_ctx* ctx, int variant);
psd_status psd_ctx_active_variant(const psd_ctx* ctx, int* variant);
typedef enum psd_constants_home { PSD_CONSTANTS_BANK = 0, PSD_CONSTANTS_SHARED = 1 } psd_constants_home;
typedef struct psd_ctx_info {
    int32_t variant;           /* what psd_ctx_active_variant reports */
    int32_t constants_home;    /* psd_constants_home of the image that kernel reads */
    uint32_t image_bytes;      /* size of that image */
    uint32_t image_shared_by;  /* live contexts on this device using the same image (>= 1) */
    char kernel_name[96];      /* e.g. "sponge_kernel<bn254,t=3,alpha=5,variant=5>" */
} psd_ctx_info;
/* No counterpart in the reference (PoseidonConfig is plain data there); lets a service see which kernel
 * and which home of constants each of its parameter sets got. */
psd_status psd_ctx_describe(const psd_ctx* ctx, psd_ctx_info* out);

#ifdef __cplusplus
}
#endif
#endif /* POSEIDON_B200_H */
