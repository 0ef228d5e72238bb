// Host (CPU) collectives over POSIX shared memory: the "gloo" role of the reference
// (test_init.py:55 `backend='gloo'`, chosen when no GPU is present, test_init.py:84-88) for
// the single-node plumbing configuration (BASELINE.json config 1: allreduce_toy, world 2,
// CPU).  One segment per process group:
//
//   [ header | result area (SLOT bytes) | rank 0 slot | rank 1 slot | ... ]
//
// all_reduce: every rank copies a chunk into its slot, barrier, rank r reduces the r-th
// 1/n of the chunk across all slots into the result area (a reduce-scatter), barrier,
// everyone copies the result out (an all-gather) - the same two-shot shape the GPU kernel
// uses over NVLink (csrc/allreduce.cu), so the CPU path exercises the same bookkeeping.
// Barriers are sense-reversing counters with a wall-clock bound: a dead peer is reported
// as an error with the ranks still missing, never a hang.
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#define B200H_API extern "C" __attribute__((visibility("default")))

namespace {

thread_local char g_err2[1024];
int fail2(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err2, sizeof(g_err2), fmt, ap);
  va_end(ap);
  return code;
}

struct Header {
  std::atomic<uint32_t> magic;
  std::atomic<uint32_t> arrived;      // barrier counter
  std::atomic<uint32_t> generation;   // barrier sense
  std::atomic<uint32_t> attached;     // ranks that mapped the segment
  std::atomic<uint64_t> heartbeat[64];
};
constexpr uint32_t kMagic = 0xB200D9u;
constexpr size_t kHeaderBytes = 4096;

struct Group {
  std::string name;
  int rank = 0, world = 1;
  size_t slot = 0, total = 0;
  uint8_t* base = nullptr;
  Header* hdr = nullptr;
  int timeout_ms = 600000;
  uint8_t* result() const { return base + kHeaderBytes; }
  uint8_t* slot_of(int r) const { return base + kHeaderBytes + slot * (size_t)(1 + r); }
};

int barrier(Group* g) {
  if (g->world == 1) return 0;
  const uint32_t gen = g->hdr->generation.load(std::memory_order_acquire);
  g->hdr->heartbeat[g->rank % 64].fetch_add(1, std::memory_order_relaxed);
  if (g->hdr->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)g->world) {
    g->hdr->arrived.store(0, std::memory_order_relaxed);
    g->hdr->generation.fetch_add(1, std::memory_order_release);
    return 0;
  }
  const auto t0 = std::chrono::steady_clock::now();
  uint32_t spins = 0;
  while (g->hdr->generation.load(std::memory_order_acquire) == gen) {
    if (++spins < 2000) continue;
    sched_yield();
    if ((spins & 0x3FF) == 0) {
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
      if (ms > g->timeout_ms)
        return fail2(1, "host barrier on group '%s' timed out after %d ms: rank %d saw only %u of %d ranks arrive",
                     g->name.c_str(), g->timeout_ms, g->rank, g->hdr->arrived.load(), g->world);
    }
  }
  return 0;
}

enum DType { F32 = 0, F64 = 1, I32 = 2, I64 = 3, U8 = 4, BF16 = 5 };
enum RedOp { SUM = 0, AVG = 1, MAX = 2, MIN = 3, PROD = 4 };

size_t dtype_size(int dt) {
  switch (dt) { case F32: case I32: return 4; case F64: case I64: return 8; case U8: return 1; case BF16: return 2; }
  return 0;
}
inline float bf16_to_f(uint16_t v) { uint32_t u = (uint32_t)v << 16; float f; memcpy(&f, &u, 4); return f; }
inline uint16_t f_to_bf16(float f) {
  uint32_t u; memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);   // NaN
  u += 0x7fffu + ((u >> 16) & 1);                                              // round to nearest even
  return (uint16_t)(u >> 16);
}

template <typename T>
void reduce_range(Group* g, size_t lo, size_t hi, int op) {
  T* out = reinterpret_cast<T*>(g->result());
  for (size_t i = lo; i < hi; ++i) {
    T acc = reinterpret_cast<const T*>(g->slot_of(0))[i];
    for (int r = 1; r < g->world; ++r) {
      const T v = reinterpret_cast<const T*>(g->slot_of(r))[i];
      switch (op) {
        case SUM: case AVG: acc = acc + v; break;
        case MAX: acc = v > acc ? v : acc; break;
        case MIN: acc = v < acc ? v : acc; break;
        case PROD: acc = acc * v; break;
      }
    }
    if (op == AVG) acc = (T)(acc / (T)g->world);
    out[i] = acc;
  }
}
void reduce_range_bf16(Group* g, size_t lo, size_t hi, int op) {
  uint16_t* out = reinterpret_cast<uint16_t*>(g->result());
  for (size_t i = lo; i < hi; ++i) {
    float acc = bf16_to_f(reinterpret_cast<const uint16_t*>(g->slot_of(0))[i]);
    for (int r = 1; r < g->world; ++r) {
      const float v = bf16_to_f(reinterpret_cast<const uint16_t*>(g->slot_of(r))[i]);
      switch (op) {
        case SUM: case AVG: acc += v; break;
        case MAX: acc = v > acc ? v : acc; break;
        case MIN: acc = v < acc ? v : acc; break;
        case PROD: acc *= v; break;
      }
    }
    if (op == AVG) acc /= (float)g->world;
    out[i] = f_to_bf16(acc);
  }
}

}  // namespace

B200H_API const char* b200h_coll_last_error() { return g_err2; }

// Rank `creator` (the lowest rank of the group) creates + initialises the segment; everyone
// else opens it (the name is published through the store by the Python layer first).
B200H_API void* b200h_coll_open(const char* name, int rank, int world, int is_creator, size_t slot_bytes,
                                int timeout_ms) {
  if (world < 1 || rank < 0 || rank >= world) { fail2(-1, "bad rank/world %d/%d", rank, world); return nullptr; }
  Group* g = new Group();
  g->name = name; g->rank = rank; g->world = world; g->slot = slot_bytes; g->timeout_ms = timeout_ms;
  g->total = kHeaderBytes + slot_bytes * (size_t)(world + 1);
  int fd = -1;
  if (is_creator) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) { fail2(-1, "shm_open(create %s): %s", name, strerror(errno)); delete g; return nullptr; }
    if (ftruncate(fd, (off_t)g->total) != 0) {
      fail2(-1, "ftruncate(%s, %zu): %s", name, g->total, strerror(errno));
      close(fd); shm_unlink(name); delete g; return nullptr;
    }
  } else {
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      fd = shm_open(name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && (size_t)st.st_size >= g->total) break;
        close(fd); fd = -1;
      }
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
      if (ms > timeout_ms) { fail2(-1, "shm_open(%s): segment did not appear within %d ms", name, timeout_ms); delete g; return nullptr; }
      usleep(1000);
    }
  }
  void* p = mmap(nullptr, g->total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) { fail2(-1, "mmap(%s): %s", name, strerror(errno)); delete g; return nullptr; }
  g->base = static_cast<uint8_t*>(p);
  g->hdr = reinterpret_cast<Header*>(p);
  if (is_creator) {
    g->hdr->arrived.store(0); g->hdr->generation.store(0); g->hdr->attached.store(0);
    g->hdr->magic.store(kMagic, std::memory_order_release);
  } else {
    const auto t0 = std::chrono::steady_clock::now();
    while (g->hdr->magic.load(std::memory_order_acquire) != kMagic) {
      auto ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
      if (ms > timeout_ms) { fail2(-1, "segment %s was never initialised by its creator", name); munmap(p, g->total); delete g; return nullptr; }
      usleep(200);
    }
  }
  g->hdr->attached.fetch_add(1);
  return g;
}

B200H_API void b200h_coll_close(void* h, int unlink_segment) {
  if (!h) return;
  Group* g = static_cast<Group*>(h);
  if (g->base) munmap(g->base, g->total);
  if (unlink_segment) shm_unlink(g->name.c_str());
  delete g;
}

B200H_API int b200h_coll_barrier(void* h) { return barrier(static_cast<Group*>(h)); }

B200H_API int b200h_coll_allreduce(void* h, void* data, size_t count, int dtype, int op) {
  Group* g = static_cast<Group*>(h);
  const size_t es = dtype_size(dtype);
  if (!es) return fail2(-2, "all_reduce: unsupported dtype code %d", dtype);
  if (op == AVG && (dtype == I32 || dtype == I64 || dtype == U8))
    return fail2(-2, "all_reduce: AVG is only defined for floating-point tensors");
  if (g->world == 1) return 0;
  const size_t per_chunk = g->slot / es;
  uint8_t* p = static_cast<uint8_t*>(data);
  for (size_t done = 0; done < count; done += per_chunk) {
    const size_t n = (count - done) < per_chunk ? (count - done) : per_chunk;
    memcpy(g->slot_of(g->rank), p + done * es, n * es);
    int rc = barrier(g);
    if (rc) return rc;
    const size_t lo = n * (size_t)g->rank / (size_t)g->world, hi = n * (size_t)(g->rank + 1) / (size_t)g->world;
    switch (dtype) {
      case F32: reduce_range<float>(g, lo, hi, op); break;
      case F64: reduce_range<double>(g, lo, hi, op); break;
      case I32: reduce_range<int32_t>(g, lo, hi, op); break;
      case I64: reduce_range<int64_t>(g, lo, hi, op); break;
      case U8: reduce_range<uint8_t>(g, lo, hi, op); break;
      case BF16: reduce_range_bf16(g, lo, hi, op); break;
    }
    rc = barrier(g);
    if (rc) return rc;
    memcpy(p + done * es, g->result(), n * es);
    rc = barrier(g);          // result area / slots are reused by the next chunk
    if (rc) return rc;
  }
  return 0;
}

B200H_API int b200h_coll_broadcast(void* h, void* data, size_t bytes, int root) {
  Group* g = static_cast<Group*>(h);
  if (root < 0 || root >= g->world) return fail2(-2, "broadcast: root %d outside group of %d", root, g->world);
  if (g->world == 1) return 0;
  uint8_t* p = static_cast<uint8_t*>(data);
  for (size_t done = 0; done < bytes; done += g->slot) {
    const size_t n = (bytes - done) < g->slot ? (bytes - done) : g->slot;
    if (g->rank == root) memcpy(g->result(), p + done, n);
    int rc = barrier(g);
    if (rc) return rc;
    if (g->rank != root) memcpy(p + done, g->result(), n);
    rc = barrier(g);
    if (rc) return rc;
  }
  return 0;
}

// out must hold world * bytes; bytes per rank must fit one slot per chunk.
B200H_API int b200h_coll_allgather(void* h, const void* in, void* out, size_t bytes) {
  Group* g = static_cast<Group*>(h);
  const uint8_t* src = static_cast<const uint8_t*>(in);
  uint8_t* dst = static_cast<uint8_t*>(out);
  if (g->world == 1) { memcpy(dst, src, bytes); return 0; }
  for (size_t done = 0; done < bytes; done += g->slot) {
    const size_t n = (bytes - done) < g->slot ? (bytes - done) : g->slot;
    memcpy(g->slot_of(g->rank), src + done, n);
    int rc = barrier(g);
    if (rc) return rc;
    for (int r = 0; r < g->world; ++r) memcpy(dst + (size_t)r * bytes + done, g->slot_of(r), n);
    rc = barrier(g);
    if (rc) return rc;
  }
  return 0;
}
