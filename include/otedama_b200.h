/*
 * otedama_b200.h — C ABI of libotedama_b200.so: a B200 (sm_100a) SHA-256d
 * proof-of-work search engine that drops in behind Otedama's
 * internal/miner worker / internal/hal device interface.
 *
 * The reference (shizukutanaka/Otedama) is pure Go and has no FFI today; every
 * entry point below names the Go symbol a cgo shim would route to it (paths are
 * relative to the reference tree).  INTEGRATION.md shows that shim.
 *
 * Conventions: plain C types only; opaque handles; caller-allocated outputs;
 * every function returning int yields 0 (or a count) on success and a negative
 * OTD_E* code on failure, with a per-thread message behind otd_last_error().
 * Hash/target byte order is the reference's miner.Hash: 32 bytes, little-endian
 * integer, most significant byte at index 31 (internal/miner/sha256d.go:84-103).
 * There is no CPU fallback: without a usable CUDA device the compute entry
 * points fail with OTD_ENODEV.
 */
#ifndef OTEDAMA_B200_H
#define OTEDAMA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OTD_OK 0
#define OTD_EINVAL (-22)
#define OTD_ENODEV (-19)
#define OTD_ENOMEM (-12)
#define OTD_EBUSY (-16)
#define OTD_ECUDA (-5)
#define OTD_ERANGE (-34)

#define OTD_ABI_VERSION 2 /* 2: job_id_str in otd_work_t / otd_share_t, job_id in otd_v1_submit_t,
                             shared_launches in otd_stats_t */

/* Last error message of the calling thread ("" if none). Never NULL. */
const char* otd_last_error(void);
int otd_abi_version(void);
/* sizeof() of the public structs as this library was compiled, so a binding can
 * check its own layout: 0 otd_device_info_t, 1 otd_header_t, 2 otd_work_t,
 * 3 otd_share_t, 4 otd_worker_config_t, 5 otd_stats_t, 6 otd_telemetry_t; -1 otherwise. */
int otd_abi_struct_size(int which);

/* ---------------------------------------------------------------- devices --
 * Replaces hal.Driver.Enumerate / hal.Device.Identity / Capabilities
 * (internal/hal/device.go:155-198) for a new "cuda" driver registered next to
 * GPULinuxDriver in engine.detectDevices (internal/engine/setup.go:29-54). */
typedef struct {
    char id[32];          /* hal.Identity.ID, e.g. "gpu-cuda0" (no whitespace or '/') */
    char model[128];      /* hal.Identity.Model */
    char vendor[16];      /* hal.Identity.Vendor: "nvidia" */
    char pci_bus_id[32];
    int32_t sm_count;
    int32_t cc_major, cc_minor;
    int32_t sm_clock_khz;
    uint64_t total_mem_bytes;
    int32_t cap_sha256d;          /* hal.Capabilities.SHA256d        (1 on sm_100) */
    int32_t cap_general_compute;  /* hal.Capabilities.GeneralCompute (1) */
} otd_device_info_t;

int otd_device_count(void);                               /* >= 0, or OTD_E* */
int otd_device_info(int device_index, otd_device_info_t* out);

/* Live device telemetry for the engine's efficiency gauges
 * (otedama_joules_per_terahash, internal/engine/metrics.go:488; per-device sampling
 * in internal/engine/stats.go:123-137): board power, SM clock, temperature, read
 * through NVML (loaded at run time; OTD_ENODEV if the driver library is absent).
 * joules per terahash = power_w / (hash_rate / 1e12). */
typedef struct {
    double power_w;
    double sm_clock_mhz;
    double temperature_c;
} otd_telemetry_t;
int otd_device_telemetry(int device_index, otd_telemetry_t* out);

/* ------------------------------------------------------------ header math --
 * Host-side mirrors of internal/miner/sha256d.go helpers (once per job). */
typedef struct {            /* miner.Header, sha256d.go:49-56 */
    uint32_t version;
    uint8_t prev_hash[32];
    uint8_t merkle_root[32];
    uint32_t time;
    uint32_t bits;
    uint32_t nonce;
} otd_header_t;

void otd_header_bytes(const otd_header_t* h, uint8_t out[80]);      /* Header.Bytes   :59-68  */
void otd_parse_header(const uint8_t in[80], otd_header_t* h);       /* ParseHeader    :71-80  */
int otd_hash_less_or_equal(const uint8_t h[32], const uint8_t other[32]); /* LessOrEqual :93-103 */
int otd_target_from_nbits(uint32_t nbits, uint8_t target[32]);      /* TargetFromNBits :134-173 */
uint32_t otd_nbits_from_target(const uint8_t target[32]);           /* NBitsFromTarget :178-205 */
int otd_target_from_difficulty(double difficulty, uint8_t target[32]); /* TargetFromDifficulty :226-247 */
/* MeetsTarget :251-257 — returns 1/0, or OTD_EINVAL for malformed nbits */
int otd_meets_target(const uint8_t hash[32], uint32_t nbits);

/* GPU batch form of miner.HashHeader / miner.SHA256d on 80-byte inputs
 * (sha256d.go:107-117): hashes n serialised headers (host pointers) on
 * `device_index`, writes n*32 bytes.  Chunks of 2^18 headers are pipelined over
 * two streams through device buffers that are allocated on the first call for a
 * device and kept.  Throughput is bound by the host-side copies of the caller's
 * pageable buffers, not by the kernel; one call at a time per device. */
int otd_hash_headers(int device_index, const uint8_t* headers80, size_t n, uint8_t* hashes32);

/* ------------------------------------------------------------------- work --
 * miner.Work (internal/miner/worker.go:18-24) plus the search-space extension
 * north_star asks for.  A zero-initialised extension reproduces the reference:
 * one header, nonce-only search of all 2^32 nonces. */
#define OTD_JOB_ID_MAX 64   /* bytes of an opaque pool job id incl. the terminating NUL */
#define OTD_WORK_REPEAT 1u /* when the space is exhausted start over (the reference's
                              uint32 wrap, worker.go:279) instead of going idle */
#define OTD_WORK_PER_VERSION_LAUNCH 2u /* rolled versions: one plain scan launch per version
                              instance instead of the version-shared kernel (A/B checks) */

typedef struct {
    uint32_t job_id;      /* Work.JobID     */
    uint32_t channel_id;  /* Work.ChannelID */
    otd_header_t header;  /* Work.Header (nonce ignored) */
    uint32_t nbits;       /* Work.NBits     */
    uint8_t target[32];   /* Work.Target — the SHARE target */

    uint32_t flags;
    /* nonce window: [nonce_begin, nonce_begin + nonce_count); nonce_count 0 = 2^32 */
    uint32_t nonce_begin;
    uint64_t nonce_count;
    /* BIP-320 version rolling: version_mask 0 = off.  Rolled version index
     * k = version_begin + i*version_stride, i in [0, version_count);
     * version_count 0 with a mask = all 2^popcount(mask) indices from
     * version_begin; stride 0 = 1.  (Static partition over G GPUs: begin = g,
     * stride = G.) */
    uint32_t version_mask;
    uint32_t version_begin, version_count, version_stride;
    /* Stratum V1 coinbase: coinb1 NULL = off (header.merkle_root is used).
     * coinbase = coinb1 || extranonce1 || extranonce2 || coinb2, extranonce2 =
     * little-endian counter of extranonce2_size (<= 16; bytes past the 8th are zero) bytes, value
     * e = extranonce2_begin + i*extranonce2_stride, i in [0, extranonce2_count).
     * merkle_branches = n_branches * 32 bytes.  Buffers are copied. */
    const uint8_t* coinb1;
    uint32_t coinb1_len;
    const uint8_t* extranonce1;
    uint32_t extranonce1_len;
    const uint8_t* coinb2;
    uint32_t coinb2_len;
    const uint8_t* merkle_branches;
    uint32_t n_branches;
    uint32_t extranonce2_size;
    uint64_t extranonce2_begin, extranonce2_count, extranonce2_stride;
    /* Stratum V1: the pool's job id exactly as mining.notify gave it (an opaque string, usually
     * hex; NUL-terminated; "" when the job only has the numeric id above).  The library never
     * parses it: it is echoed into every share of this work so that mining.submit can send it
     * back verbatim.  (The reference squeezes the id through Sscanf("%d") into a uint32,
     * internal/engine/run.go:1345-1348, which only works for decimal ids.) */
    char job_id_str[OTD_JOB_ID_MAX];
} otd_work_t;

typedef struct {          /* miner.Share, worker.go:33-43 */
    uint32_t channel_id;
    uint32_t job_id;
    uint32_t nonce;
    uint32_t ntime;
    uint32_t version;     /* the exact (rolled) version that was hashed */
    uint8_t hash[32];
    uint64_t extranonce2; /* value rolled into the coinbase (0 when unused) */
    uint64_t work_seq;    /* which set_work call this share belongs to (1-based) */
    int32_t device_index; /* -> Share.DeviceID via otd_device_info().id */
    int32_t reserved;
    char job_id_str[OTD_JOB_ID_MAX]; /* the work's job_id_str, echoed */
} otd_share_t;

/* ------------------------------------------------------- job -> work glue --
 * The engine-side functions that turn a pool job into a miner.Work
 * (internal/engine/run.go); pure host code, no device needed.
 *
 * otd_v1_job_target — v1JobTarget (run.go:1313-1324): the nBits block target,
 * replaced by the pool's share-difficulty target when difficulty > 0 and that
 * conversion succeeds.  Malformed nbits fail regardless of difficulty.
 * otd_work_from_v2_job — updateWork (run.go:1271-1297): all five header inputs
 * from a Stratum V2 standard-channel job + chain tip; the share target, or the
 * nBits target when the pool assigned none (all-zero).  Returns OTD_EINVAL (and
 * leaves *work untouched) where the reference silently returns.
 * otd_work_from_v1_job — applyJob (run.go:1340-1364) as the reference has it:
 * job id parsed like Sscanf("%d"), header carries merkle root / time / bits only
 * (version and prev-hash are NOT copied by the reference). */
int otd_v1_job_target(uint32_t nbits, double difficulty, uint8_t target[32]);
int otd_work_from_v2_job(uint32_t job_id, uint32_t channel_id, uint32_t version, const uint8_t merkle_root[32],
                         const uint8_t prev_hash[32], uint32_t prev_nbits, uint32_t ntime,
                         const uint8_t share_target[32], otd_work_t* work);
int otd_work_from_v1_job(const char* job_id, uint32_t channel_id, const uint8_t merkle_root[32], uint32_t ntime,
                         uint32_t nbits, double difficulty, otd_work_t* work);

/* Stratum V1 mining.notify -> device work unit with on-device extranonce2 rolling
 * (north_star).  The reference's parseNotify keeps job id, prevhash,
 * version, nbits, ntime and DISCARDS coinb1 / coinb2 / merkle_branch
 * (internal/poolproto/stratumv1/parse.go:33-93), so this follows the public
 * Stratum V1 conventions ("parity unpinned" against the reference; pinned instead by
 * the protocol documentation's own worked example and by block 170):
 *   params = [job_id, prevhash, coinb1, coinb2, merkle_branch[], version, nbits, ntime, clean_jobs]
 *   - prevhash: 64 hex chars, eight 4-byte words each byte-swapped with respect to
 *     header order (header bytes = each word reversed);
 *   - version / nbits / ntime: big-endian hex of the u32 header fields;
 *   - coinbase = coinb1 || extranonce1 || extranonce2 || coinb2 (hex decoded).
 * job_id stays an opaque string (work->job_id_str, at most OTD_JOB_ID_MAX - 1 characters,
 * echoed into the shares); work->job_id holds its value only when the whole string is a
 * decimal number that fits 32 bits, else 0 — Stratum job ids are usually hex, which the
 * reference's applyJob (Sscanf "%d", run.go:1345-1348; otd_work_from_v1_job mirrors it) would
 * reject or truncate.  The share target comes from v1JobTarget(nbits, difficulty).  The decoded bytes
 * the work points to are placed in `scratch` (caller-owned; must outlive the use
 * of *work; otd_worker_set_work / otd_scan copy them).  Returns OTD_OK, or
 * OTD_ERANGE with *scratch_needed set when scratch_len is too small. */
typedef struct {
    const char* job_id;
    const char* prevhash_hex;
    const char* coinb1_hex;
    const char* coinb2_hex;
    const char* const* merkle_branch_hex; /* n_branches strings of 64 hex chars */
    uint32_t n_branches;
    const char* version_hex;
    const char* nbits_hex;
    const char* ntime_hex;
    int32_t clean_jobs;
} otd_v1_notify_t;
int otd_work_from_v1_notify(const otd_v1_notify_t* notify, const char* extranonce1_hex,
                            uint32_t extranonce2_size, double difficulty, uint32_t channel_id,
                            otd_work_t* work, uint8_t* scratch, size_t scratch_len, size_t* scratch_needed);

/* A found share -> the hex fields of Stratum V1 mining.submit
 * (["worker", job_id, extranonce2, ntime, nonce] and, with version rolling, the
 * BIP-310 sixth field version_bits).
 *   job_id           = share.job_id_str verbatim, or "%u" of share.job_id when that is empty;  Replaces the body of session.Submit's
 * parameter formatting (internal/poolproto/stratumv1/stratumv1.go:336-346),
 * which today pads extranonce2 with zeros because the reference never rolls it:
 *   extranonce2_hex  = the extranonce2_size bytes that were placed in the coinbase
 *                      (little-endian counter, like otd_work_t says), lower-case hex;
 *   ntime_hex, nonce_hex = "%08x" of the u32 values, as the reference formats them;
 *   version_bits_hex = "%08x" of share.version & version_mask ("" when the mask is 0).
 * extranonce2_size 0 (no coinbase in the work) gives an empty extranonce2_hex.
 * OTD_EINVAL for a NULL argument or extranonce2_size > 16. */
typedef struct {
    char extranonce2_hex[33];
    char ntime_hex[9];
    char nonce_hex[9];
    char version_bits_hex[9];
    char job_id[OTD_JOB_ID_MAX];
} otd_v1_submit_t;
int otd_share_to_v1_submit(const otd_share_t* share, uint32_t extranonce2_size, uint32_t version_mask,
                           otd_v1_submit_t* out);

/* Static, zero-communication partition of a work's search space over G devices
 * (the reference's only parallelism is nonce striding inside one
 * Worker, internal/miner/worker.go:51-60,279).  Device g of G gets, in order of
 * preference: every G-th BIP-320 version index (version_mask set), else every
 * G-th extranonce2 value (coinbase set), else the g-th of G contiguous nonce
 * windows.  *out is a shallow copy of *in with the slice fields rewritten (it
 * points at the same coinbase buffers).  OTD_EINVAL when g is out of range or
 * the axis has fewer than G values. */
int otd_work_partition(const otd_work_t* in, uint32_t g, uint32_t G, otd_work_t* out);

/* ----------------------------------------------------------------- worker --
 * One worker per device: the GPU counterpart of *miner.Worker
 * (internal/miner/worker.go:91-212), constructed by engine.startMinerWorkers
 * (internal/engine/setup.go:59-77) for every device with SHA256d capability. */
typedef struct otd_worker otd_worker_t;

typedef struct {
    uint32_t ring_capacity;   /* device share ring entries per launch; 0 = 4096 */
    uint32_t queue_capacity;  /* host share queue (reference: chan cap Threads*4); 0 = 65536.
                                 Shares arriving at a full queue are dropped and counted. */
    uint32_t launch_log2;     /* candidates per kernel launch = 2^launch_log2 (16..32); 0 = 32.  A version-shared
                                 launch of g instances covers 2^launch_log2 / g nonces (otd_plan_launches) */
    int32_t kernel_variant;   /* scan kernel variant, -1 = library default (see otd_scan_variants) */
} otd_worker_config_t;

typedef struct {            /* miner.Stats, worker.go:79-86, plus device extras */
    uint64_t hashes_total;
    uint64_t shares_found;
    uint64_t shares_dropped;
    uint64_t uptime_ns;
    double hash_rate;         /* lifetime average, like the reference */
    uint64_t launches;        /* scan kernel launches completed */
    uint64_t candidates;      /* candidates that passed the top-word test (slow path runs) */
    double device_ms;         /* scan-kernel time on the device (CUDA events) */
    uint64_t work_seq;        /* current set_work generation */
    int32_t has_work;         /* Worker.HasWork() */
    int32_t exhausted;        /* 1 when the current work's search space is finished */
    uint64_t shared_launches; /* of `launches`: version-shared ones (BIP-320 rolled work, >= 2 instances each) */
} otd_stats_t;

/* NewWorker + Start (worker.go:111-151).  cfg may be NULL. Returns NULL on error. */
otd_worker_t* otd_worker_create(int device_index, const otd_worker_config_t* cfg);
/* SetWork (worker.go:171-176). work == NULL pauses (idle). The work and its
 * buffers are copied; safe from any thread. Returns the new work_seq (> 0). */
int64_t otd_worker_set_work(otd_worker_t* w, const otd_work_t* work);
/* The receive side of Start's `<-chan Share`: blocks up to timeout_ms
 * (0 = poll, < 0 = forever) until at least one share is available, the current
 * work has ENDED (its search space is exhausted, or its run failed) and no poll
 * has returned for that yet, or the worker is destroyed.  Returns the count
 * written.  The end of a work wakes exactly one empty poll: with 0 it if the
 * space was exhausted, with OTD_ECUDA (message in otd_last_error) if the run
 * failed; after that empty polls wait for their timeout again until
 * otd_worker_set_work hands over new work (which also clears the failure). */
int otd_worker_poll_shares(otd_worker_t* w, otd_share_t* out, int max_shares, int timeout_ms);
/* Blocks until the current work is exhausted (or timeout). 1 = exhausted, 0 = timeout. */
int otd_worker_wait_exhausted(otd_worker_t* w, int timeout_ms);
int otd_worker_stats(otd_worker_t* w, otd_stats_t* out);        /* Stats()   worker.go:189-212 */
int otd_worker_has_work(otd_worker_t* w);                       /* HasWork() worker.go:181-185 */
int otd_worker_device_index(otd_worker_t* w);                   /* DeviceID() worker.go:176    */
/* Stop (worker.go:156-165): aborts in-flight work, joins the host thread, frees.
 * All other calls on a handle may run concurrently with each other; destroy must
 * be the last one — a caller parked in otd_worker_poll_shares has to return
 * before the handle is destroyed (the Go shim waits for its forwarding goroutine,
 * INTEGRATION.md §2), because the handle's memory is released here. */
int otd_worker_destroy(otd_worker_t* w);

/* ------------------------------------------------------- synchronous scan --
 * One pass of grind (worker.go:216-282) over `work`'s whole search space on one
 * device, returning when done: the call tests and benchmarks use.  Writes up to
 * max_shares shares (sorted by extranonce2, version, nonce) and returns how many
 * it wrote; *n_found gets the total found on the device, which can be larger when
 * max_shares or the per-launch device ring (65536 entries) was too small.  Optional outputs may be NULL:
 * *hashes = candidates hashed, *device_ms = scan-kernel time by CUDA events on
 * the launching stream, *launches = scan kernel launches. */
int otd_scan(int device_index, const otd_work_t* work, int kernel_variant, otd_share_t* out,
             uint32_t max_shares, uint64_t* n_found, uint64_t* hashes, double* device_ms,
             uint64_t* launches);

/* Burst form for many independent headers (BASELINE configs[1]: genesis + 256
 * synthetic headers): header i of the batch is searched over the nonce window
 * against `target`; every launch is queued between one pair of CUDA events with
 * no host synchronisation in between (*device_ms is that span).  Shares carry the
 * header's batch index in job_id and come back sorted by (index, nonce).  The
 * whole burst shares one device ring of 65536 entries: more shares than that are
 * dropped and show up only in *n_found.  Returns the number of shares written. */
int otd_scan_headers(int device_index, const otd_header_t* headers, uint32_t n_headers,
                     const uint8_t target[32], uint32_t nonce_begin, uint64_t nonce_count,
                     int kernel_variant, otd_share_t* out, uint32_t max_shares, uint64_t* n_found,
                     uint64_t* hashes, double* device_ms, uint64_t* launches);

/* otd_scan_noalloc — no counterpart in the reference (Go code never deals with
 * device memory).  It is otd_scan for callers that must not allocate: a host
 * process whose GPU memory is owned up front by another allocator or capped by
 * its operator, or a latency-critical caller for which cudaMalloc/cudaFree (which
 * synchronise the whole device and can fail under memory pressure) are off
 * limits.  The call allocates nothing on the device or the host heap: the
 * kernel's scratch — share ring, its counters, the stop word — is a fixed
 * OTD_NOALLOC_SCRATCH_BYTES array inside the library that every GPU loading it
 * holds until the process exits.
 *   - Plain-header work only (no version rolling, no coinbase); the nonce window
 *     may be anything up to 2^32.
 *   - max_shares may not exceed OTD_NOALLOC_MAX_SHARES (the ring's size): a larger
 *     request fails with OTD_ERANGE.  More shares than max_shares found on the
 *     device are dropped and show up in *n_found.
 *   - The first call on a device looks up the array's device address; make that
 *     first call before starting any CUDA stream capture.
 *   - All calls share the one array and it keeps the previous call's contents;
 *     each call clears the counters it needs.  The call is synchronous and the
 *     library lets only one run at a time — two scans using the array at once on
 *     the GPU would race, so do not try to overlap them from outside either
 *     (another process gets its own copy of the array).
 *   - The clear, the kernel and the copies run on the legacy default stream
 *     (stream 0): they order themselves after, and hold back, work the host
 *     application has queued on its own blocking streams of that device; streams
 *     created with cudaStreamNonBlocking are not affected.
 * Output is the same as otd_scan's for the same work (sorted by nonce) as long as
 * nothing was dropped.  When the device finds more than max_shares, the ones kept
 * are whichever reached the ring first — an arbitrary subset that can differ from
 * run to run (otd_scan keeps the lowest of the full set); *n_found tells.  Returns
 * the number of shares written. */
#define OTD_NOALLOC_MAX_SHARES 4096u
#define OTD_NOALLOC_SCRATCH_BYTES (256u + 40u * OTD_NOALLOC_MAX_SHARES) /* 164,096 bytes per GPU */
int otd_scan_noalloc(int device_index, const otd_work_t* work, int kernel_variant, otd_share_t* out,
                     uint32_t max_shares, uint64_t* n_found, uint64_t* hashes);

/* Kernel variants compiled into the library: returns the count, fills
 * names[i] (each up to 32 bytes incl. NUL) when names != NULL. */
int otd_scan_variants(char (*names)[32], int max_names);
int otd_scan_default_variant(void);

/* Version-shared scan variants.  A work that rolls BIP-320 version bits
 * (version_mask != 0, at least two version instances in the slice) is searched by
 * a second kernel family: the versions of a job differ only in block 1 of the
 * header, so block 2's message schedule is expanded once per nonce and reused by
 * up to 64 version instances per launch (csrc/scan_vs_kernel.cuh).  No reference
 * counterpart: the reference ignores version rolling
 * (internal/poolproto/stratumv1/stratumv1.go:280-281).  The share set is the same
 * as with one plain launch per version (OTD_WORK_PER_VERSION_LAUNCH).
 * otd_vs_select picks the variant every later run of this process uses for rolled
 * work: an index, -1 = library default, -2 = never share (plain kernel per version). */
int otd_vs_variants(char (*names)[32], int max_names);
int otd_vs_default_variant(void);
int otd_vs_select(int variant);

/* The launch plan of a work: how one pass over its search space is cut into kernel
 * launches — (first instance, instances in the launch, nonce window) — exactly as
 * otd_scan and the worker do it.  Pure host arithmetic, no device needed; exposed so
 * that the partition of the space can be property-tested on a CPU box (every
 * (instance, nonce) in exactly one launch; a version-shared launch holds at most
 * OTD_VS_MAX_INSTANCES instances of ONE extranonce2 value).  No reference counterpart
 * (one goroutine strides one header's nonces, internal/miner/worker.go:248-280).
 * launch_log2 as in otd_worker_config_t (0 = 32).  Writes up to max_launches entries,
 * returns how many; *n_launches (may be NULL) gets the total, which walks the whole
 * plan: refused with OTD_ERANGE for a work of more than 2^22 instances. */
#define OTD_VS_MAX_INSTANCES 64u
typedef struct {
    uint64_t first_instance;      /* instance = (extranonce2 index) * versions_per_extranonce2 + version slot */
    uint32_t n_instances;         /* consecutive instances in this launch */
    uint32_t version_shared;      /* 1 = version-shared kernel, 0 = plain kernel */
    uint32_t nonce_begin;
    uint32_t version_index_first; /* BIP-320 index of the first instance (0 without a mask) */
    uint64_t nonce_count;
    uint64_t extranonce2;         /* value of the launch's instances (0 without a coinbase) */
} otd_launch_plan_t;
int otd_plan_launches(const otd_work_t* work, uint32_t launch_log2, otd_launch_plan_t* out,
                      uint32_t max_launches, uint64_t* n_launches);

/* Job-preparation kernels exposed for parity tests (K3: extranonce2 roll +
 * coinbase hash + merkle fold; K2: rolled version). Fills n merkle roots
 * (32 bytes each) / n versions for the work's first n instances. */
int otd_prepare_instances(int device_index, const otd_work_t* work, uint32_t first, uint32_t n,
                          uint8_t* merkle_roots, uint32_t* versions, uint64_t* extranonce2s,
                          uint8_t* midstates /* n*32 bytes, BE words as bytes; may be NULL */);

/* ---------------------------------------------------- integer-issue peak --
 * Measures the sustained integer instruction rate of the device (the roofline
 * this path is bound by).  mode: see otd_int_peak_modes.  Returns thread-level
 * integer ops per second in *ops_per_s (instructions of the named class per
 * thread, summed over the chip) and the SM clock seen (*sm_mhz, from clock64
 * ticks over event time). */
int otd_int_peak(int device_index, int mode, int iters, double* ops_per_s, double* sm_mhz,
                 double* ms);
int otd_int_peak_modes(char (*names)[32], int max_names);

#ifdef __cplusplus
}
#endif
#endif /* OTEDAMA_B200_H */
