// fa100_ptx.cuh -- thin inline-PTX wrappers for sm_100a: mbarrier, TMA, tcgen05 (MMA / TMEM).
// Encodings cross-checked against cute/arch/{mma_sm100_desc,mma_sm100_umma,copy_sm100,
// tmem_allocator_sm100}.hpp (CUTLASS 4.x headers) -- nothing from CUTLASS is included or linked.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace fa100 {

#define FA_DEVICE __device__ __forceinline__

// ------------------------------------------------------------------ misc
FA_DEVICE uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

FA_DEVICE bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

template <int N> FA_DEVICE void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(N)); }
template <int N> FA_DEVICE void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(N)); }

FA_DEVICE void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory"); }

// ------------------------------------------------------------------ programmatic dependent launch (PDL)
// Role of cutlass::arch::wait_on_dependent_grids / launch_dependent_grids in the reference (hopper/flash_fwd_kernel_sm90.h,
// flash_fwd_combine_kernel.h, flash_prepare_scheduler.cu).  Contract in this library: every kernel executes pdl_wait() in
// all of its threads BEFORE its first global-memory access, so a kernel launched with the programmatic-serialization
// attribute only overlaps its prologue (barrier init, TMEM allocation, descriptor prefetch) and its launch latency with the
// previous kernel's tail, and completion of a kernel implies completion of everything before it in the stream.  Both
// instructions are no-ops in a kernel launched without the attribute.
FA_DEVICE void pdl_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
FA_DEVICE void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }

// ------------------------------------------------------------------ mbarrier
FA_DEVICE void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
FA_DEVICE void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
FA_DEVICE void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// Jitter build (-DFA100_JITTER, libfa100_jitter.so; never the product library): a pseudo-random pause of 0 .. ~2 us in
// front of every barrier wait and arrive of every role, so that the roles interleave differently in every key block.  The
// parity and bitwise-repeat tests must still pass on it: a hand-off that only works by timing breaks (our stand-in for
// compute-sanitizer racecheck, which is closed on this pool -- profiles/r02_sanitizer_closed_on_pool.log).
#ifdef FA100_JITTER
FA_DEVICE void fa_jitter() {
  uint32_t x = uint32_t(clock()) * 2654435761u + (threadIdx.x >> 5) * 40503u + blockIdx.x * 9176u;
  x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
  if ((x & 3u) == 0u) __nanosleep((x >> 8) & 2047u);  // one time in four; warp-uniform enough: lanes of a warp read nearly the same clock
}
#else
FA_DEVICE void fa_jitter() {}
#endif

FA_DEVICE void mbar_arrive(uint64_t* bar) {
  fa_jitter();
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
FA_DEVICE void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
FA_DEVICE bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// Non-blocking look at a phase (try_wait may park the thread for a while when the phase is still open).
FA_DEVICE bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// ------------------------------------------------------------------ CTA pairs (cta_group::2)
// Encodings follow cute/arch/copy_sm100_tma.hpp (SM100_TMA_2SM_LOAD), cutlass/arch/barrier.h (umma_arrive_multicast_2x1SM,
// ClusterBarrier::arrive(cta_id)) and cute/arch/tmem_allocator_sm100.hpp (Allocator2Sm).
FA_DEVICE uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r)); return r; }
FA_DEVICE void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// shared::cluster address of the same shared-memory location in CTA `rank` of the cluster
FA_DEVICE uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// Default semantics (release at CTA scope), as cutlass::arch::ClusterBarrier::arrive(cta_id) does: what the arrival orders
// here are tensor-memory stores, already fenced by tcgen05.wait::st / fence::before_thread_sync.  A `.release.cluster`
// arrive costs ~1100 clk per call (cluster-scope release = cache maintenance), measured in profiles/r01c_buckets_pair.log.
FA_DEVICE void mbar_arrive_cluster(uint32_t cluster_addr) {
  fa_jitter();
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];\n" ::"r"(cluster_addr) : "memory");
}
// the pair's leader (even CTA) owns the barriers the MMA warps wait on: clearing the peer bit of a shared::cluster
// address names the leader's copy
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;

// Watchdog: a wrong phase/count must never hang the GPU box.  A wait that lasts longer than
// FA100_WATCHDOG_CYCLES records {tag, block, thread, parity} (read by fa100_debug_hang_info()) and TRAPS: the launch
// fails with a CUDA error at the caller's next synchronisation instead of finishing with garbage.  The abort flag makes
// the other CTAs' waits fall through so that the grid drains quickly.
#ifndef FA100_WATCHDOG_CYCLES
#define FA100_WATCHDOG_CYCLES (4000000000ll)
#endif
static __device__ unsigned int g_fa100_hang_info[4];
static __device__ volatile unsigned int g_fa100_abort;

// Opt-in timeline (build with -DFA100_TRACE -> libfa100_trace.so): lane 0 of every warp of one CTA records
// {tag | warp << 16, begin, end} (SM clock, low 32 bits) for each barrier wait and for explicit FA_TRACE_POINTs;
// the host reads it with fa100_debug_trace().  Never compiled into the product library.
#ifdef FA100_TRACE
#ifndef FA100_TRACE_BLOCK
#define FA100_TRACE_BLOCK 0
#endif
constexpr int kTraceWarps = 20, kTracePerWarp = 4096;
static __device__ unsigned int g_fa100_trace[kTraceWarps * kTracePerWarp * 3];
static __device__ unsigned int g_fa100_trace_cnt[kTraceWarps];
#ifndef FA100_TRACE_MAXWARP
#define FA100_TRACE_MAXWARP 8  // events from the softmax warps only: event stores in the issuing warp distort what they measure
#endif
FA_DEVICE bool trace_on() { return blockIdx.x == FA100_TRACE_BLOCK && (threadIdx.x & 31) == 0; }
FA_DEVICE bool trace_events_on() { return trace_on() && (threadIdx.x >> 5) < FA100_TRACE_MAXWARP; }
FA_DEVICE unsigned int* trace_counter() {  // per-warp event counter in shared memory (a global counter costs an L2 round trip per event)
  // last 128 bytes of the dynamic allocation (free tail of the barrier block): a static array would push the
  // kernel over the 227 KB limit
  extern __shared__ __align__(1024) uint8_t fa100_trace_smem[];
  uint32_t dyn;
  asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
  return reinterpret_cast<unsigned int*>(fa100_trace_smem + dyn - 128) + (threadIdx.x >> 5);
}
FA_DEVICE void trace_begin() { if ((threadIdx.x & 31) == 0) *trace_counter() = 0; }
FA_DEVICE void trace_end() { if (trace_on()) g_fa100_trace_cnt[threadIdx.x >> 5] = *trace_counter(); }
FA_DEVICE void trace_emit(int tag, long long tb, long long te) {
  const int w = threadIdx.x >> 5;
  unsigned int* cnt = trace_counter();
  const unsigned int n = *cnt;
  if (n < kTracePerWarp) {
    unsigned int* dst = g_fa100_trace + (w * kTracePerWarp + n) * 3;
    dst[0] = unsigned(tag) | (unsigned(w) << 16); dst[1] = unsigned(tb); dst[2] = unsigned(te);
    *cnt = n + 1;
  }
}
#ifdef FA100_TRACE_NOEVENTS  // libfa100_prof.so: cycle buckets only, no per-event stores to perturb the timing
#define FA_TRACE_POINT(tag) do { } while (0)
#else
#define FA_TRACE_POINT(tag) do { if (trace_events_on()) { long long t_ = clock64(); trace_emit(tag, t_, t_); } } while (0)
#endif
#define FA_TRACE_BEGIN() trace_begin()
#define FA_TRACE_END() trace_end()
// Cycle accounting without memory traffic in the loop: FA_PROF(i) charges the cycles since the previous mark to
// bucket i (registers only); one CTA writes its buckets out at the end.  fa100_debug_prof() reads them.
static __device__ unsigned int g_fa100_prof[kTraceWarps * 8];
#define FA_PROF_DECL() uint32_t prof_[8] = {0, 0, 0, 0, 0, 0, 0, 0}; uint32_t prof_t_ = clock()
#define FA_PROF(i) do { const uint32_t n_ = clock(); prof_[i] += n_ - prof_t_; prof_t_ = n_; } while (0)
// latency probes: one role stamps the SM clock into shared memory, another charges "now - stamp" to a bucket
#define FA_LAT_STAMP(word) do { *reinterpret_cast<volatile uint32_t*>(word) = clock(); } while (0)
#define FA_LAT_ACC(i, word) do { prof_[i] += clock() - *reinterpret_cast<volatile uint32_t*>(word); } while (0)
#define FA_PROF_FLUSH() do { if (trace_on()) { for (int i_ = 0; i_ < 8; ++i_) g_fa100_prof[(threadIdx.x >> 5) * 8 + i_] = prof_[i_]; } } while (0)
#else
#define FA_PROF_DECL() do { } while (0)
#define FA_PROF(i) do { } while (0)
#define FA_PROF_FLUSH() do { } while (0)
#define FA_LAT_STAMP(word) do { } while (0)
#define FA_LAT_ACC(i, word) do { } while (0)
#define FA_TRACE_POINT(tag) do { } while (0)
#define FA_TRACE_BEGIN() do { } while (0)
#define FA_TRACE_END() do { } while (0)
#endif

// Slow path, kept as small as possible because it is inlined at a dozen wait sites and the kernel's three roles
// already overflow the instruction cache: poll; every 64 polls look at the abort flag and the clock.
FA_DEVICE void mbar_wait_slow(uint64_t* bar, uint32_t parity, int tag) {
  const long long t0 = clock64();
  uint32_t polls = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++polls & 63u) == 0) {
      if (g_fa100_abort) return;
      if (clock64() - t0 > FA100_WATCHDOG_CYCLES) {
        g_fa100_hang_info[0] = 0xDEAD0000u | (unsigned)tag;  // last writer wins: any stuck wait identifies the bug
        g_fa100_hang_info[1] = blockIdx.x;
        g_fa100_hang_info[2] = threadIdx.x;
        g_fa100_hang_info[3] = parity;
        g_fa100_abort = 1;
        __threadfence();
        __trap();
      }
    }
  }
}

FA_DEVICE void mbar_wait(uint64_t* bar, uint32_t parity, int tag = 0) {
  fa_jitter();
#if defined(FA100_TRACE) && !defined(FA100_TRACE_NOEVENTS)
  const long long tb_ = clock64();
#endif
  if (!mbar_try_wait(bar, parity) && !mbar_try_wait(bar, parity)) mbar_wait_slow(bar, parity, tag);
#if defined(FA100_TRACE) && !defined(FA100_TRACE_NOEVENTS)
  if (trace_events_on()) trace_emit(tag, tb_, clock64());
#endif
}

// ------------------------------------------------------------------ TMA
FA_DEVICE void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];\n" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
FA_DEVICE void tma_load_4d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];\n"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
FA_DEVICE void tma_load_5d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];\n"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
// 2-CTA loads: the data lands in this CTA's shared memory, the bytes are counted on the LEADER's mbarrier
FA_DEVICE void tma_load_4d_2sm(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];\n"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
FA_DEVICE void tma_load_5d_2sm(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];\n"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
FA_DEVICE void tma_store_4d(const CUtensorMap* m, const void* src, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];\n"
      ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
FA_DEVICE void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
FA_DEVICE void tma_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }
FA_DEVICE void tma_store_wait0() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }

// ------------------------------------------------------------------ tcgen05: TMEM management
FA_DEVICE void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
}
FA_DEVICE void tmem_relinquish() { asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory"); }
FA_DEVICE void tmem_dealloc(uint32_t addr, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(addr), "r"(ncols) : "memory");
}
FA_DEVICE void tmem_alloc_2sm(uint32_t* smem_dst, uint32_t ncols) {  // the same warp of both CTAs of a pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
}
FA_DEVICE void tmem_relinquish_2sm() { asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory"); }
FA_DEVICE void tmem_dealloc_2sm(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(addr), "r"(ncols) : "memory");
}
FA_DEVICE void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
FA_DEVICE void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

// tcgen05.commit: the mbarrier gets one arrival once all MMAs issued so far by this thread are done.
FA_DEVICE void tc_commit(uint64_t* bar) {
  fa_jitter();
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}

// 2-CTA commit: one arrival on the barrier at the same offset in BOTH CTAs of the pair once the MMAs issued so far are done
FA_DEVICE void tc_commit_2sm(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
               ::"r"(smem_u32(bar)), "h"(uint16_t(3)) : "memory");
}

// ------------------------------------------------------------------ tcgen05: MMA
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16 and fp16 inputs with fp32 accumulate.
FA_DEVICE void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// A operand read from TMEM (row i of A in lane i; 16-bit elements packed two per 32-bit column).
FA_DEVICE void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
FA_DEVICE void mma_f8_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
FA_DEVICE void mma_f8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Same instructions with the 64-bit shared-memory descriptors given as {lo, hi} words: hi is a compile-time
// constant and lo = (address >> 4) + offset | LBO field, so the issue loop spends one integer add per MMA.
template <bool FP8>
FA_DEVICE void mma_ss(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate) {
  if (FP8)
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], da, db, %4, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
template <bool FP8>
FA_DEVICE void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate) {
  if (FP8)
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], db, %4, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
  else
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n\t}\n"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// cta_group::2 forms (16-bit operands): M = 256 rows over the pair, every CTA supplies its 128 rows of A and half of B
FA_DEVICE void mma_ss_2sm(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
FA_DEVICE void mma_ts_2sm(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.b32 p, %5, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], db, %4, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// descriptor words for a 128-byte-swizzled operand: hi = SBO >> 4 | version 1 << 14 | SWIZZLE_128B << 29
__host__ __device__ constexpr uint32_t sdesc_hi(uint32_t sbo_bytes) { return ((sbo_bytes >> 4) & 0x3FFF) | (1u << 14) | (2u << 29); }
__host__ __device__ constexpr uint32_t sdesc_lo_lbo(uint32_t lbo_bytes) { return ((lbo_bytes >> 4) & 0x3FFF) << 16; }

// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format[4,6) a_format[7,10) b_format[10,13)
// a_major[15] b_major[16] (0 = K-major, 1 = MN-major) n_dim[17,23) = N>>3, m_dim[24,29) = M>>4.
enum : uint32_t { kFmtF16 = 0, kFmtBF16 = 1, kFmtE4M3 = 0, kFmtE5M2 = 1 };
__host__ __device__ constexpr uint32_t make_idesc(uint32_t a_fmt, uint32_t b_fmt, uint32_t M, uint32_t N,
                                                   uint32_t a_mn_major, uint32_t b_mn_major) {
  return (1u << 4) | (a_fmt << 7) | (b_fmt << 10) | (a_mn_major << 15) | (b_mn_major << 16) | ((N >> 3) << 17) |
         ((M >> 4) << 24);
}

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor), 128-byte swizzle, version 1:
// start>>4 [0,14)  LBO>>4 [16,30)  SBO>>4 [32,46)  version [46,48)=1  layout_type [61,64)=2 (SWIZZLE_128B).
__host__ __device__ constexpr uint64_t make_sdesc_base(uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t((lbo_bytes >> 4) & 0x3FFF) << 16) | (uint64_t((sbo_bytes >> 4) & 0x3FFF) << 32) | (uint64_t(1) << 46) |
         (uint64_t(2) << 61);
}
FA_DEVICE uint64_t sdesc_at(uint64_t base, uint32_t smem_addr) { return base | uint64_t((smem_addr & 0x3FFFF) >> 4); }

// ------------------------------------------------------------------ tcgen05: TMEM <-> registers
// 32x32b: thread l of warp w (w % 4 selects the 32-lane quarter) reads lane 32*(w%4)+l, N consecutive columns.
FA_DEVICE void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
FA_DEVICE void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

FA_DEVICE void tmem_ld32(uint32_t addr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(addr)
      : "memory");
}
FA_DEVICE void tmem_st32(uint32_t addr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};\n"
      ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
      "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
      "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
FA_DEVICE void tmem_st16(uint32_t addr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n"
      ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
FA_DEVICE void tmem_st8(uint32_t addr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n"
      ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
      : "memory");
}

}  // namespace fa100
