// Single-node low-latency all-reduce between host processes through one POSIX shared-memory segment.
//
// Layout of the segment:  Header | slot[world][chunk_bytes] | result[chunk_bytes]
// One chunk goes through four phases; two process-shared sense-reversing barriers separate them:
//   1. every rank copies its chunk into slot[rank]
//   2. barrier; rank r sums element slice r of ALL slots (fp32 accumulation) into result  -- the reduction is split
//      over the ranks (a reduce-scatter into `result`), no rank is a serial bottleneck
//   3. barrier; every rank copies the whole result out
// The next chunk's phase 1 only touches the rank's own slot, and nobody can reach the next phase 2 before every rank
// has left phase 3 (barrier of the next chunk), so two barriers per chunk suffice.
// Reference counterpart: csrc/cpu/comm/ccl.cpp:70-96,146-272,540-611 (`inference_all_reduce`: state-machine flags, rank 0
// reduces everything).  Used by comm/ for host-resident tensor-parallel inference (`inference_all_reduce`).
#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>

#define DSBH_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

constexpr uint32_t kMagic = 0x44534248;  // "DSBH"

struct alignas(64) Header {
  std::atomic<uint32_t> magic;
  int world;
  long long chunk_bytes;
  alignas(64) std::atomic<int> arrived;
  alignas(64) std::atomic<int> sense;
  alignas(64) std::atomic<int> attached;
};

struct Handle {
  Header* hdr;
  uint8_t* slots;
  uint8_t* result;
  size_t map_bytes;
  int rank, world;
  long long chunk_bytes;
  int local_sense;
  char name[128];
};

inline void cpu_relax() {
#if defined(__x86_64__)
  __builtin_ia32_pause();
#endif
}

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

// returns 0, or -1 after `timeout_s` without every rank arriving (a dead peer must not hang the survivors forever)
int barrier(Handle* h, double timeout_s) {
  Header* H = h->hdr;
  h->local_sense ^= 1;
  if (H->arrived.fetch_add(1, std::memory_order_acq_rel) == h->world - 1) {
    H->arrived.store(0, std::memory_order_relaxed);
    H->sense.store(h->local_sense, std::memory_order_release);
    return 0;
  }
  const double t0 = now_s();
  uint32_t spins = 0;
  while (H->sense.load(std::memory_order_acquire) != h->local_sense) {
    cpu_relax();
    if ((++spins & 0xfff) == 0) {
      if (now_s() - t0 > timeout_s) return -1;
      if (spins > (1u << 20)) usleep(50);  // oversubscribed hosts: stop burning the core the peer needs
    }
  }
  return 0;
}

inline float bf16_to_f32(uint16_t v) {
  uint32_t u = static_cast<uint32_t>(v) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(u >> 16);
}

// dtype: 0 = fp32, 1 = bf16, 2 = fp16 (F16C), 3 = int32 ; sums elements [lo, hi) of every slot into result
void reduce_slice(const Handle* h, int dtype, long long lo, long long hi) {
  const int W = h->world;
  const long long stride = h->chunk_bytes;
  if (dtype == 0) {
    float* out = reinterpret_cast<float*>(h->result);
    for (long long i = lo; i < hi; ++i) {
      float acc = 0.f;
      for (int r = 0; r < W; ++r) acc += reinterpret_cast<const float*>(h->slots + r * stride)[i];
      out[i] = acc;
    }
  } else if (dtype == 1) {
    uint16_t* out = reinterpret_cast<uint16_t*>(h->result);
    for (long long i = lo; i < hi; ++i) {
      float acc = 0.f;
      for (int r = 0; r < W; ++r) acc += bf16_to_f32(reinterpret_cast<const uint16_t*>(h->slots + r * stride)[i]);
      out[i] = f32_to_bf16(acc);
    }
  } else if (dtype == 2) {
    _Float16* out = reinterpret_cast<_Float16*>(h->result);
    for (long long i = lo; i < hi; ++i) {
      float acc = 0.f;
      for (int r = 0; r < W; ++r) acc += static_cast<float>(reinterpret_cast<const _Float16*>(h->slots + r * stride)[i]);
      out[i] = static_cast<_Float16>(acc);
    }
  } else {
    int32_t* out = reinterpret_cast<int32_t*>(h->result);
    for (long long i = lo; i < hi; ++i) {
      int32_t acc = 0;
      for (int r = 0; r < W; ++r) acc += reinterpret_cast<const int32_t*>(h->slots + r * stride)[i];
      out[i] = acc;
    }
  }
}

int elem_size(int dtype) { return (dtype == 0 || dtype == 3) ? 4 : 2; }

}  // namespace

// Creates (rank 0) or attaches to the segment `/name`.  Returns an opaque handle or nullptr.
DSBH_EXPORT void* dsbh_shm_open(const char* name, int rank, int world, long long chunk_bytes, double timeout_s) {
  if (!name || world < 1 || rank < 0 || rank >= world || chunk_bytes < 64 || strlen(name) > 100) return nullptr;
  chunk_bytes = (chunk_bytes + 63) & ~63LL;
  const size_t hdr_bytes = (sizeof(Header) + 4095) & ~size_t(4095);
  const size_t total = hdr_bytes + static_cast<size_t>(world + 1) * chunk_bytes;
  char path[128];
  snprintf(path, sizeof(path), "/%s", name);
  int fd = -1;
  const double t0 = now_s();
  if (rank == 0) {
    shm_unlink(path);  // a stale segment of a crashed job
    fd = shm_open(path, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return nullptr;
    if (ftruncate(fd, static_cast<off_t>(total)) != 0) {
      close(fd);
      shm_unlink(path);
      return nullptr;
    }
  } else {
    while (true) {
      fd = shm_open(path, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && static_cast<size_t>(st.st_size) >= total) break;
        close(fd);
        fd = -1;
      }
      if (now_s() - t0 > timeout_s) return nullptr;
      usleep(1000);
    }
  }
  void* base = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (base == MAP_FAILED) return nullptr;
  Header* H = reinterpret_cast<Header*>(base);
  if (rank == 0) {
    H->world = world;
    H->chunk_bytes = chunk_bytes;
    H->arrived.store(0);
    H->sense.store(0);
    H->attached.store(0);
    H->magic.store(kMagic, std::memory_order_release);
  } else {
    while (H->magic.load(std::memory_order_acquire) != kMagic) {
      if (now_s() - t0 > timeout_s) {
        munmap(base, total);
        return nullptr;
      }
      usleep(200);
    }
    if (H->world != world || H->chunk_bytes != chunk_bytes) {
      munmap(base, total);
      return nullptr;
    }
  }
  Handle* h = new Handle();
  h->hdr = H;
  h->slots = reinterpret_cast<uint8_t*>(base) + hdr_bytes;
  h->result = h->slots + static_cast<size_t>(world) * chunk_bytes;
  h->map_bytes = total;
  h->rank = rank;
  h->world = world;
  h->chunk_bytes = chunk_bytes;
  h->local_sense = 0;
  snprintf(h->name, sizeof(h->name), "%s", path);
  H->attached.fetch_add(1);
  // everybody attached before the creator may unlink the name
  while (H->attached.load(std::memory_order_acquire) < world) {
    if (now_s() - t0 > timeout_s) break;
    usleep(200);
  }
  if (rank == 0) shm_unlink(path);  // the mapping stays alive; nothing is left behind if the job dies
  return h;
}

// In-place sum over all ranks of `numel` elements at `data`.  Returns 0, 1 (bad argument) or 2 (peer timeout).
DSBH_EXPORT int dsbh_shm_allreduce(void* handle, void* data, long long numel, int dtype, double timeout_s) {
  Handle* h = reinterpret_cast<Handle*>(handle);
  if (!h || !data || numel < 0 || dtype < 0 || dtype > 3) return 1;
  const int esz = elem_size(dtype);
  const long long per_chunk = h->chunk_bytes / esz;
  uint8_t* p = reinterpret_cast<uint8_t*>(data);
  for (long long done = 0; done < numel || (numel == 0 && done == 0); done += per_chunk) {
    const long long n = numel - done < per_chunk ? numel - done : per_chunk;
    if (n > 0) memcpy(h->slots + static_cast<size_t>(h->rank) * h->chunk_bytes, p + done * esz, static_cast<size_t>(n) * esz);
    if (barrier(h, timeout_s)) return 2;
    // slice boundaries on 16-element multiples: no two ranks write the same cache line of `result`
    const long long per_rank = (((n + h->world - 1) / h->world) + 15) & ~15LL;
    const long long lo = per_rank * h->rank < n ? per_rank * h->rank : n;
    const long long hi = lo + per_rank < n ? lo + per_rank : n;
    if (hi > lo) reduce_slice(h, dtype, lo, hi);
    if (barrier(h, timeout_s)) return 2;
    if (n > 0) memcpy(p + done * esz, h->result, static_cast<size_t>(n) * esz);
    if (numel == 0) break;
  }
  return 0;
}

DSBH_EXPORT int dsbh_shm_barrier(void* handle, double timeout_s) {
  Handle* h = reinterpret_cast<Handle*>(handle);
  if (!h) return 1;
  return barrier(h, timeout_s) ? 2 : 0;
}

DSBH_EXPORT void dsbh_shm_close(void* handle) {
  Handle* h = reinterpret_cast<Handle*>(handle);
  if (!h) return;
  munmap(reinterpret_cast<uint8_t*>(h->hdr), h->map_bytes);
  delete h;
}
