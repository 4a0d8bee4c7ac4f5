// Native epoch executor + data feeder: the hot loop of reference src/trainer.py:180-197
// (DataLoader.__next__ -> .to(device) -> step -> loss.item()) without the Python interpreter in it.
//
// Per step the executor does, entirely from C++:
//   gather thread : collect the step's rows of an in-memory (host) dataset into a pinned staging slot
//   copy stream   : cudaMemcpyAsync H2D of inputs + targets into the slot's device buffers, event `ready`
//   compute stream: wait `ready`; H2D of the step's 8-float hyper-parameter row (lr schedule precomputed by the
//                   caller); launch the step (either the single-kernel LeNet step or a captured CUDA graph whose
//                   input pointers are the slot's buffers); D2H of the step's loss into a pinned result array;
//                   event `released` (the slot's buffers may be overwritten).
// Slots form a ring (default 3), so the gather and upload of step i+1.. overlap the compute of step i.
// Nothing here synchronises with the device except when a ring slot is still in flight and at the end of the run.
//
// Fixed cost per run is kept to a few tens of microseconds: the gather thread lives as long as the loop (parked on a
// condition variable between runs; the caller gathers step 0 itself while it wakes), the per-run tables are allocated once,
// and a slot's inputs and targets share ONE pinned buffer and ONE device buffer, so a full batch goes device-ward as one copy.
#include <cuda_runtime.h>
#include <emmintrin.h>
#include <linux/futex.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "../include/mlt_b200.h"

namespace mlt {
void set_error(const char* fmt, ...);
}
using mlt::set_error;

#define RT_CHECK(expr)                                                                    \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));     \
      return (int)_e;                                                                     \
    }                                                                                     \
  } while (0)

namespace {

struct Slot {
  void* pin_x = nullptr;            // pinned staging: [max_batch x rows | pad to 256 B | max_batch y rows]
  void* pin_y = nullptr;            // = pin_x + y_off
  void* dev_x = nullptr;            // device buffer, same layout
  void* dev_y = nullptr;            // = dev_x + y_off
  cudaEvent_t ready = nullptr;      // H2D of this slot finished (copy stream)
  cudaEvent_t copied = nullptr;     // same event viewed from the host: pinned staging reusable
  cudaEvent_t released = nullptr;   // compute finished with the device buffers
  bool used = false;
};

struct Loop {
  int device = 0;
  int slots = 3;
  long long max_batch = 0;
  size_t x_row = 0, y_row = 0;          // bytes per sample
  std::vector<Slot> ring;
  cudaStream_t copy = nullptr;
  float* pin_hyper = nullptr;           // [max_steps][8] staged hyper rows
  long long hyper_cap = 0;
  float* pin_loss = nullptr;            // [max_steps]
  long long loss_cap = 0;
  float* dev_hyper = nullptr;           // [max_steps][8] device copy of the whole epoch's hyper rows (kernel steps)
  long long dev_hyper_cap = 0;
  // persistent LeNet runs (step kind 2): the whole epoch is ONE kernel launch that waits for each slot to be filled
  unsigned int* dev_ready = nullptr;    // device word the copy stream advances to i + 1 after step i's batch has landed
  unsigned int* pin_progress = nullptr; // pinned host word the kernel advances to i + 1 after step i
  unsigned int* pin_seq = nullptr;      // pinned [max_steps] = 1, 2, 3, ... (source of the ready updates)
  long long seq_cap = 0;
  int* dev_bs = nullptr;                // [max_steps] per-step batch sizes
  long long bs_cap = 0;
  // statistics of the last run
  long long h2d_bytes = 0, d2h_bytes = 0, launches = 0;
  // Large samples (images): skip the host gather of x and let the copy engine gather instead - one cudaMemcpyAsync per
  // sample straight from the (page-locked) dataset into the slot's device buffer.
  bool direct_x = false;
  size_t y_off = 0;                     // byte offset of the targets inside a slot's buffers
  // uint8 image datasets: augment on the device before every step (augment.cu)
  bool aug_on = false;
  mlt_augment_cfg aug{};
  void* aug_out = nullptr;
  int aug_bf16 = 0, aug_train = 1;
  unsigned long long aug_pos0 = 0;
  // ---- gather thread (one per loop, parked between runs) ------------------------------------------------------
  struct Job {
    const char* hx = nullptr; const char* hy = nullptr;
    const long long* indices = nullptr; const long long* start = nullptr;
    long long first = 0, steps = 0;
  } job;
  // kMaxWorkers gather threads at most (MLT_GATHER_THREADS, default 2): worker w fills the steps first + w, first + w + W, ...
  static constexpr int kMaxWorkers = 4;
  int nworkers = 1;
  std::thread worker[kMaxWorkers];
  std::mutex mu;
  std::condition_variable cv;
  long long job_seq = 0;                // guarded by mu
  bool stop = false;                    // guarded by mu
  std::atomic<long long> done_seq[kMaxWorkers];   // last job each worker finished
  std::atomic<bool> abort{false};
  // Waiting without burning a core: the two counters below are bumped (and futex-woken) whenever `gathered` / `free_for`
  // change; a waiter spins briefly, then sleeps on the counter.  Eight ranks x three threads that all spin do not fit a box that
  // gives each rank a handful of cores.  MLT_WAIT_SPIN=1 restores pure spinning.
  std::atomic<uint32_t> w_gathered{0}, w_free{0};
  bool pure_spin = false;
  std::unique_ptr<std::atomic<long long>[]> gathered;   // [slots] step id (+1) whose data sits in the slot's pinned buffers
  std::unique_ptr<std::atomic<long long>[]> free_for;   // [slots] slot may be refilled for any step <= this value
  std::vector<long long> start;         // prefix sums of the per-step batch sizes of the current run
  // host-side timing of the last run, microseconds since entry: [first launch enqueued, all launches enqueued, device drained];
  // [3] = time the enqueueing thread spent waiting for gathered batches
  double t_us[4] = {0, 0, 0, 0};
};

void futex_wake_all(std::atomic<uint32_t>& word) {
  word.fetch_add(1, std::memory_order_release);
  syscall(SYS_futex, reinterpret_cast<uint32_t*>(&word), FUTEX_WAKE_PRIVATE, 0x7fffffff, nullptr, nullptr, 0);
}

// Block until ready() holds; `word` is bumped by whoever may make it true (after the state change).
template <typename Pred>
void wait_until(std::atomic<uint32_t>& word, bool pure_spin, Pred ready) {
  for (int spin = 0; spin < 256; ++spin) {
    if (ready()) return;
    _mm_pause();
  }
  for (;;) {
    const uint32_t seen = word.load(std::memory_order_acquire);
    if (ready()) return;
    if (pure_spin) { std::this_thread::yield(); continue; }
    timespec ts{0, 500000};               // wake up at least every 0.5 ms (abort flags are polled by the predicates)
    syscall(SYS_futex, reinterpret_cast<uint32_t*>(&word), FUTEX_WAIT_PRIVATE, seen, &ts, nullptr, 0);
  }
}

// Copy into a pinned staging buffer.  The CPU never reads the destination (the copy engine does), so large rows are written with
// non-temporal stores: no read-for-ownership of the destination lines and no cache pollution.  The caller publishes the slot
// with a release store AFTER the sfence below.
void stage_copy(char* dst, const char* src, size_t bytes) {
  if (bytes < 2048 || ((uintptr_t)dst & 15)) { std::memcpy(dst, src, bytes); return; }
  const size_t blocks = bytes / 64;
  const __m128i* s = (const __m128i*)src;
  __m128i* d = (__m128i*)dst;
  for (size_t b = 0; b < blocks; ++b) {
    const __m128i v0 = _mm_loadu_si128(s + 4 * b), v1 = _mm_loadu_si128(s + 4 * b + 1);
    const __m128i v2 = _mm_loadu_si128(s + 4 * b + 2), v3 = _mm_loadu_si128(s + 4 * b + 3);
    _mm_stream_si128(d + 4 * b, v0); _mm_stream_si128(d + 4 * b + 1, v1);
    _mm_stream_si128(d + 4 * b + 2, v2); _mm_stream_si128(d + 4 * b + 3, v3);
  }
  if (bytes % 64) std::memcpy(dst + blocks * 64, src + blocks * 64, bytes % 64);
}

// rows `idx[0..n)` of a [N][row] byte matrix -> dst
void gather_rows(char* dst, const char* src, size_t row, const long long* idx, long long n) {
  bool contiguous = true;
  for (long long i = 1; i < n; ++i)
    if (idx[i] != idx[i - 1] + 1) { contiguous = false; break; }
  if (contiguous) stage_copy(dst, src + (size_t)idx[0] * row, (size_t)n * row);
  else
    for (long long i = 0; i < n; ++i) stage_copy(dst + (size_t)i * row, src + (size_t)idx[i] * row, row);
  _mm_sfence();
}

void gather_step(Loop* L, const Loop::Job& j, long long i) {
  const int s = (int)(i % L->slots);
  const long long n = j.start[i + 1] - j.start[i];
  if (!L->direct_x) gather_rows((char*)L->ring[s].pin_x, j.hx, L->x_row, j.indices + j.start[i], n);
  gather_rows((char*)L->ring[s].pin_y, j.hy, L->y_row, j.indices + j.start[i], n);
  L->gathered[s].store(i + 1, std::memory_order_release);
  futex_wake_all(L->w_gathered);
}

void worker_main(Loop* L, int w) {
  long long seen = 0;
  for (;;) {
    Loop::Job j;
    {
      std::unique_lock<std::mutex> lk(L->mu);
      L->cv.wait(lk, [&] { return L->stop || L->job_seq != seen; });
      if (L->stop) return;
      seen = L->job_seq;
      j = L->job;
    }
    for (long long i = j.first + w; i < j.steps && !L->abort.load(std::memory_order_relaxed); i += L->nworkers) {
      const int s = (int)(i % L->slots);
      wait_until(L->w_free, L->pure_spin, [&] {
        return L->free_for[s].load(std::memory_order_acquire) >= i || L->abort.load(std::memory_order_relaxed);
      });
      if (L->abort.load(std::memory_order_relaxed)) break;
      gather_step(L, j, i);
    }
    L->done_seq[w].store(seen, std::memory_order_release);
  }
}

double now_us() {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

extern "C" void* mlt_loop_create(int device, int slots, long long max_batch, long long x_row_bytes, long long y_row_bytes) {
  if (cudaSetDevice(device) != cudaSuccess) { set_error("mlt_loop_create: cudaSetDevice(%d) failed", device); return nullptr; }
  Loop* L = new Loop();
  L->device = device;
  L->slots = slots < 2 ? 2 : slots;
  L->max_batch = max_batch;
  L->x_row = (size_t)x_row_bytes;
  L->y_row = (size_t)y_row_bytes;
  L->ring.resize(L->slots);
  bool ok = cudaStreamCreateWithFlags(&L->copy, cudaStreamNonBlocking) == cudaSuccess;
  L->y_off = ((size_t)max_batch * L->x_row + 255) / 256 * 256;
  const size_t slot_bytes = L->y_off + (size_t)max_batch * L->y_row;
  for (auto& s : L->ring) {
    ok = ok && cudaHostAlloc(&s.pin_x, slot_bytes, cudaHostAllocDefault) == cudaSuccess;
    ok = ok && cudaMalloc(&s.dev_x, slot_bytes) == cudaSuccess;
    if (ok) { s.pin_y = (char*)s.pin_x + L->y_off; s.dev_y = (char*)s.dev_x + L->y_off; }
    ok = ok && cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&s.released, cudaEventDisableTiming) == cudaSuccess;
  }
  // per-run tables, sized for 8192 steps up front (they grow on demand)
  const long long cap = 8192;
  ok = ok && cudaHostAlloc((void**)&L->pin_hyper, (size_t)cap * 8 * sizeof(float), cudaHostAllocDefault) == cudaSuccess;
  ok = ok && cudaHostAlloc((void**)&L->pin_loss, (size_t)cap * sizeof(float), cudaHostAllocDefault) == cudaSuccess;
  ok = ok && cudaMalloc((void**)&L->dev_hyper, (size_t)cap * 8 * sizeof(float)) == cudaSuccess;
  if (!ok) {
    set_error("mlt_loop_create: allocation failed (%s)", cudaGetErrorString(cudaGetLastError()));
    delete L;
    return nullptr;
  }
  L->hyper_cap = L->loss_cap = L->dev_hyper_cap = cap;
  L->gathered.reset(new std::atomic<long long>[L->slots]);
  L->free_for.reset(new std::atomic<long long>[L->slots]);
  const char* ws = std::getenv("MLT_WAIT_SPIN");
  L->pure_spin = ws && std::atoi(ws) == 1;
  const char* gt = std::getenv("MLT_GATHER_THREADS");
  L->nworkers = gt ? std::atoi(gt) : 2;
  if (L->nworkers < 1) L->nworkers = 1;
  if (L->nworkers > Loop::kMaxWorkers) L->nworkers = Loop::kMaxWorkers;
  if (L->nworkers > L->slots - 1) L->nworkers = L->slots - 1;        // a worker per slot that is not being consumed
  for (int w = 0; w < L->nworkers; ++w) { L->done_seq[w].store(0); L->worker[w] = std::thread(worker_main, L, w); }
  return L;
}

extern "C" void* mlt_loop_slot_x(void* loop, int slot) { return ((Loop*)loop)->ring[slot].dev_x; }
extern "C" void* mlt_loop_slot_y(void* loop, int slot) { return ((Loop*)loop)->ring[slot].dev_y; }
extern "C" int mlt_loop_num_slots(void* loop) { return ((Loop*)loop)->slots; }
// on != 0: the dataset's x tensor is page-locked; copy samples device-ward one by one instead of gathering on the host
extern "C" int mlt_loop_set_direct_rows(void* loop, int on) { ((Loop*)loop)->direct_x = on != 0; return 0; }

extern "C" int mlt_loop_set_augment(void* loop, const mlt_augment_cfg* cfg, void* out_dev, int out_bf16, unsigned long long pos0,
                                    int train) {
  Loop* L = (Loop*)loop;
  L->aug_on = cfg != nullptr && out_dev != nullptr;
  if (L->aug_on) { L->aug = *cfg; L->aug_out = out_dev; L->aug_bf16 = out_bf16; L->aug_pos0 = pos0; L->aug_train = train; }
  return 0;
}

extern "C" int mlt_loop_destroy(void* loop) {
  Loop* L = (Loop*)loop;
  if (!L) return 0;
  { std::lock_guard<std::mutex> lk(L->mu); L->stop = true; }
  L->cv.notify_all();
  for (int w = 0; w < L->nworkers; ++w)
    if (L->worker[w].joinable()) L->worker[w].join();
  cudaSetDevice(L->device);
  cudaDeviceSynchronize();
  for (auto& s : L->ring) {
    cudaFreeHost(s.pin_x); cudaFree(s.dev_x);
    cudaEventDestroy(s.ready); cudaEventDestroy(s.released);
  }
  if (L->pin_hyper) cudaFreeHost(L->pin_hyper);
  if (L->pin_loss) cudaFreeHost(L->pin_loss);
  if (L->dev_hyper) cudaFree(L->dev_hyper);
  if (L->dev_ready) cudaFree(L->dev_ready);
  if (L->pin_progress) cudaFreeHost(L->pin_progress);
  if (L->pin_seq) cudaFreeHost(L->pin_seq);
  if (L->dev_bs) cudaFree(L->dev_bs);
  cudaStreamDestroy(L->copy);
  delete L;
  return 0;
}

// Run `steps` training (or evaluation) steps.
//   host_x / host_y   in-memory dataset, row-major, x_row_bytes / y_row_bytes per sample (pageable or pinned)
//   indices           [steps * batch] sample ids, step-major (the sampler's order for this rank)
//   batch_sizes       per-step batch size (<= max_batch); NULL = every step uses `batch`
//   hyper_rows        [steps][8] floats copied to `hyper_dev` before each step (NULL for evaluation)
//   step kind 0       LeNet single-kernel step: lenet points at the fixed arguments of mlt_lenet_step
//   step kind 2       the same LeNet step as ONE persistent launch for the whole run (mlt_lenet_steps): the kernel waits on a
//                     device word the copy stream advances after each batch, the host watches a pinned progress word
//   step kind 1       captured CUDA graphs, one cudaGraphExec_t per slot in graph_execs (inputs bound to the slot buffers)
//   loss_dev          device scalar the step writes its loss to; copied to loss_host[step] every step
extern "C" int mlt_loop_run(void* loop, const void* host_x, const void* host_y, const long long* indices, long long steps,
                            int batch, const int* batch_sizes, const float* hyper_rows, void* hyper_dev, int step_kind,
                            const mlt_lenet_call* lenet, void* const* graph_execs, const void* loss_dev, float* loss_host,
                            void* compute_stream) {
  Loop* L = (Loop*)loop;
  const double t_enter = now_us();
  L->t_us[3] = 0;
  RT_CHECK(cudaSetDevice(L->device));
  cudaStream_t cs = (cudaStream_t)compute_stream;
  if (steps <= 0) return 0;
  if (batch > L->max_batch) { set_error("mlt_loop_run: batch %d exceeds the loop's max_batch %lld", batch, L->max_batch); return 1; }
  // per-run tables are sized generously on first use: a pinned (re)allocation costs a few hundred microseconds, which would
  // otherwise land inside the first longer epoch
  const long long cap = steps < 8192 ? 8192 : steps + steps / 2;
  if (hyper_rows && steps > L->hyper_cap) {
    if (L->pin_hyper) cudaFreeHost(L->pin_hyper);
    RT_CHECK(cudaHostAlloc((void**)&L->pin_hyper, (size_t)cap * 8 * sizeof(float), cudaHostAllocDefault));
    L->hyper_cap = cap;
  }
  if (hyper_rows) std::memcpy(L->pin_hyper, hyper_rows, (size_t)steps * 8 * sizeof(float));
  if (loss_dev && steps > L->loss_cap) {
    if (L->pin_loss) cudaFreeHost(L->pin_loss);
    RT_CHECK(cudaHostAlloc((void**)&L->pin_loss, (size_t)cap * sizeof(float), cudaHostAllocDefault));
    L->loss_cap = cap;
  }
  L->h2d_bytes = L->d2h_bytes = L->launches = 0;
  // Single-kernel steps read their hyper-parameter row straight from a device table uploaded once per run, and write
  // their loss straight into the pinned host array (a posted store over PCIe): no per-step copy operations at all.
  const bool table_mode = step_kind == 0 || step_kind == 2;
  const bool persistent = step_kind == 2;
  if (persistent && L->aug_on) {
    set_error("mlt_loop_run: the persistent step kernel owns every SM; it cannot be combined with the per-step augmentation kernel");
    return 1;
  }
  if (table_mode && hyper_rows) {
    if (steps > L->dev_hyper_cap) {
      if (L->dev_hyper) cudaFree(L->dev_hyper);
      RT_CHECK(cudaMalloc((void**)&L->dev_hyper, (size_t)cap * 8 * sizeof(float)));
      L->dev_hyper_cap = cap;
    }
    RT_CHECK(cudaMemcpyAsync(L->dev_hyper, L->pin_hyper, (size_t)steps * 8 * sizeof(float), cudaMemcpyHostToDevice, cs));
    L->h2d_bytes += steps * 32;
  }

  // ---- gather: the caller fills the pinned slot of step 0 itself while the parked gather thread wakes up for steps 1.. ----
  const int S = L->slots;
  for (int s = 0; s < S; ++s) { L->gathered[s].store(0); L->free_for[s].store(s); }
  L->abort.store(false);
  const char* hx = (const char*)host_x;
  const char* hy = (const char*)host_y;
  auto bs_of = [&](long long i) { return batch_sizes ? (long long)batch_sizes[i] : (long long)batch; };
  std::vector<long long>& start = L->start;
  start.assign(steps + 1, 0);
  for (long long i = 0; i < steps; ++i) start[i + 1] = start[i] + bs_of(i);
  long long my_seq;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->job.hx = hx; L->job.hy = hy; L->job.indices = indices; L->job.start = start.data();
    L->job.first = 1; L->job.steps = steps;
    my_seq = ++L->job_seq;
  }
  L->cv.notify_all();
  gather_step(L, L->job, 0);
  std::atomic<bool>& abort = L->abort;
  auto& gathered = L->gathered;
  auto& free_for = L->free_for;
  struct Joiner {            // every exit path waits until the gather thread is done with this run's pointers
    Loop* L; long long seq;
    void join() {
      futex_wake_all(L->w_free);           // a worker asleep on a slot (abort paths) sees the flag now
      for (int w = 0; w < L->nworkers; ++w)
        while (L->done_seq[w].load(std::memory_order_acquire) != seq) std::this_thread::yield();
    }
  } gatherer{L, my_seq};

  int rc = 0;
  if (persistent) {
    // ONE launch for the whole run: the kernel walks the ring, gated by *dev_ready, and reports through *pin_progress
    auto fail = [&](const char* what) { abort.store(true); gatherer.join(); set_error("mlt_loop_run: %s", what); return 1; };
    if (S > 8 || steps > 0x7fffffffLL) return fail("persistent runs need <= 8 slots");
    if (!L->dev_ready && cudaMalloc((void**)&L->dev_ready, 256) != cudaSuccess) return fail("cudaMalloc failed");
    if (!L->pin_progress && cudaHostAlloc((void**)&L->pin_progress, 256, cudaHostAllocDefault) != cudaSuccess) return fail("cudaHostAlloc failed");
    if (steps > L->seq_cap) {
      if (L->pin_seq) cudaFreeHost(L->pin_seq);
      if (cudaHostAlloc((void**)&L->pin_seq, (size_t)steps * 4, cudaHostAllocDefault) != cudaSuccess) return fail("cudaHostAlloc failed");
      for (long long i = 0; i < steps; ++i) L->pin_seq[i] = (unsigned int)(i + 1);
      L->seq_cap = steps;
    }
    if (batch_sizes && steps > L->bs_cap) {
      if (L->dev_bs) cudaFree(L->dev_bs);
      if (cudaMalloc((void**)&L->dev_bs, (size_t)steps * 4) != cudaSuccess) return fail("cudaMalloc failed");
      L->bs_cap = steps;
    }
    if (batch_sizes && cudaMemcpyAsync(L->dev_bs, batch_sizes, (size_t)steps * 4, cudaMemcpyHostToDevice, cs) != cudaSuccess) return fail("batch-size upload failed");
    if (cudaMemsetAsync(L->dev_ready, 0, 4, cs) != cudaSuccess) return fail("memset failed");
    *(volatile unsigned int*)L->pin_progress = 0u;
    const void* xs[8]; const void* ys[8];
    for (int k = 0; k < S; ++k) { xs[k] = L->ring[k].dev_x; ys[k] = L->ring[k].dev_y; }
    if (cudaStreamSynchronize(cs) != cudaSuccess) return fail("stream sync failed");      // tables and the zeroed ready word are in place
    if (mlt_lenet_steps(lenet->master, lenet->s1, lenet->s2, lenet->hyper, lenet->numel, lenet->param_off, lenet->rule, xs, ys, S, 0,
                        batch, batch_sizes ? L->dev_bs : nullptr, hyper_rows ? (const void*)L->dev_hyper : nullptr,
                        loss_dev ? (void*)L->pin_loss : nullptr, L->dev_ready, L->pin_progress, (int)steps, lenet->train,
                        lenet->want_acc, lenet->acc3, lenet->scratch, lenet->scratch_bytes, lenet->ctx, cs) != 0) {
      abort.store(true); gatherer.join(); return 1;
    }
    L->launches += 1;
  }
  for (long long i = 0; i < steps; ++i) {
    const int s = (int)(i % S);
    Slot& sl = L->ring[s];
    const long long n = bs_of(i);
    // Hand the previous step's pinned staging back to the gather thread once its H2D has completed (it has had a
    // whole iteration to do so; block only if it has not).
    if (i > 0 && i - 1 + S < steps) {
      Slot& pv = L->ring[(i - 1) % S];
      if (cudaEventQuery(pv.ready) != cudaSuccess) cudaEventSynchronize(pv.ready);
      free_for[(i - 1) % S].store(i - 1 + S, std::memory_order_release);
      futex_wake_all(L->w_free);
    }
    if (gathered[s].load(std::memory_order_acquire) != i + 1) {
      const double tw = now_us();
      wait_until(L->w_gathered, L->pure_spin, [&] { return gathered[s].load(std::memory_order_acquire) == i + 1; });
      L->t_us[3] += now_us() - tw;
    }
    // device buffers of this slot are free once the step that last used them has finished
    if (persistent) {
      if (i >= S) {
        const unsigned int need = (unsigned int)(i - S + 1);
        volatile unsigned int* prog = L->pin_progress;
        long long spins = 0;
        while (*prog < need) {
          if (*prog == 0xFFFFFFFFu || ++spins > (1LL << 34)) { rc = (int)cudaErrorLaunchTimeout; break; }
          if ((spins & 1023) == 0) std::this_thread::yield();
        }
        if (rc != 0) break;
      }
    } else if (sl.used) { if ((rc = (int)cudaStreamWaitEvent(L->copy, sl.released, 0)) != 0) break; }
    if (L->direct_x) {
      const long long* idx = indices + start[i];
      for (long long j = 0; j < n && rc == 0; ) {
        long long run = 1;                                  // coalesce runs of consecutive sample ids into one copy
        while (j + run < n && idx[j + run] == idx[j + run - 1] + 1) ++run;
        rc = (int)cudaMemcpyAsync((char*)sl.dev_x + (size_t)j * L->x_row, hx + (size_t)idx[j] * L->x_row, (size_t)run * L->x_row,
                                  cudaMemcpyHostToDevice, L->copy);
        j += run;
      }
      if (rc != 0) break;
    } else if (n == L->max_batch) {
      // full batch: inputs and targets are adjacent in the slot, ONE copy carries both
      if ((rc = (int)cudaMemcpyAsync(sl.dev_x, sl.pin_x, L->y_off + (size_t)n * L->y_row, cudaMemcpyHostToDevice, L->copy)) != 0) break;
    } else {
      if ((rc = (int)cudaMemcpyAsync(sl.dev_x, sl.pin_x, (size_t)n * L->x_row, cudaMemcpyHostToDevice, L->copy)) != 0) break;
    }
    if (L->direct_x || n != L->max_batch)
      if ((rc = (int)cudaMemcpyAsync(sl.dev_y, sl.pin_y, (size_t)n * L->y_row, cudaMemcpyHostToDevice, L->copy)) != 0) break;
    if ((rc = (int)cudaEventRecord(sl.ready, L->copy)) != 0) break;
    L->h2d_bytes += n * (long long)(L->x_row + L->y_row);
    if (persistent) {
      // publish "step i has landed" behind the copies on the same stream; the running kernel polls the word
      if ((rc = (int)cudaMemcpyAsync(L->dev_ready, L->pin_seq + i, 4, cudaMemcpyHostToDevice, L->copy)) != 0) break;
      if (loss_dev) L->d2h_bytes += 4;
      continue;
    }
    if ((rc = (int)cudaStreamWaitEvent(cs, sl.ready, 0)) != 0) break;
    if (L->aug_on) {      // uint8 rows -> cropped / flipped / normalised batch the step reads
      if ((rc = mlt_augment_u8(sl.dev_x, L->aug_out, L->aug_bf16, (int)n, &L->aug, L->aug_pos0 + (unsigned long long)start[i], L->aug_train, cs)) != 0) {
        abort.store(true); gatherer.join(); return rc;
      }
      L->launches += 1;
    }
    if (hyper_rows && !table_mode) {
      if ((rc = (int)cudaMemcpyAsync(hyper_dev, L->pin_hyper + i * 8, 8 * sizeof(float), cudaMemcpyHostToDevice, cs)) != 0) break;
      L->h2d_bytes += 32;
    }
    if (step_kind == 0) {
      const void* hy = hyper_rows ? (const void*)(L->dev_hyper + i * 8) : lenet->hyper;
      void* lo = loss_dev ? (void*)(L->pin_loss + i) : lenet->loss_out;      // pinned host memory is device-addressable (UVA)
      rc = mlt_lenet_step(lenet->master, lenet->s1, lenet->s2, hy, lenet->numel, lenet->param_off, lenet->rule,
                          L->aug_on ? L->aug_out : sl.dev_x, sl.dev_y, (int)n, lenet->train, lenet->want_acc, lenet->acc3, lo,
                          lenet->scratch, lenet->scratch_bytes, lenet->ctx, cs);
      if (rc != 0) break;
      if (loss_dev) L->d2h_bytes += 4;
    } else {
      if ((rc = (int)cudaGraphLaunch((cudaGraphExec_t)graph_execs[s], cs)) != 0) break;
    }
    L->launches += 1;
    if (i == 0) L->t_us[0] = now_us() - t_enter;
    if (loss_dev && !table_mode) {
      if ((rc = (int)cudaMemcpyAsync(L->pin_loss + i, loss_dev, sizeof(float), cudaMemcpyDeviceToHost, cs)) != 0) break;
      L->d2h_bytes += 4;
    }
    if ((rc = (int)cudaEventRecord(sl.released, cs)) != 0) break;
    sl.used = true;
  }
  if (rc != 0) {
    abort.store(true);
    set_error("mlt_loop_run: CUDA call failed with %s", cudaGetErrorString((cudaError_t)rc));
  }
  gatherer.join();
  if (rc != 0) return rc;
  L->t_us[1] = now_us() - t_enter;
  RT_CHECK(cudaStreamSynchronize(cs));
  L->t_us[2] = now_us() - t_enter;
  if (persistent && *(volatile unsigned int*)L->pin_progress != (unsigned int)steps) {
    set_error("mlt_loop_run: the persistent step kernel stopped after %u of %lld steps", *(volatile unsigned int*)L->pin_progress, steps);
    return 1;
  }
  if (loss_dev && loss_host) std::memcpy(loss_host, L->pin_loss, (size_t)steps * sizeof(float));
  return 0;
}

// Host-side timing of the last run in microseconds since entry: out3 = {first step enqueued, all steps enqueued, device drained}.
// mlt_loop_gather_wait_us: how long the enqueueing thread waited for gathered batches in that run.
extern "C" double mlt_loop_gather_wait_us(void* loop) { return ((Loop*)loop)->t_us[3]; }
extern "C" int mlt_loop_timing(void* loop, double* out3) {
  Loop* L = (Loop*)loop;
  for (int i = 0; i < 3; ++i) out3[i] = L->t_us[i];
  return 0;
}

extern "C" int mlt_loop_stats(void* loop, long long* h2d_bytes, long long* d2h_bytes, long long* launches) {
  Loop* L = (Loop*)loop;
  if (h2d_bytes) *h2d_bytes = L->h2d_bytes;
  if (d2h_bytes) *d2h_bytes = L->d2h_bytes;
  if (launches) *launches = L->launches;
  return 0;
}
