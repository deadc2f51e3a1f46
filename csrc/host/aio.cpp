// Asynchronous file I/O for the NVMe tier of ZeRO-Infinity: a handle with a worker-thread pool that splits each
// request into block_size chunks and issues positional reads/writes (O_DIRECT when the buffer and sizes allow it),
// sync and async variants with wait(), as the reference's aio_handle
// (csrc/aio/py_lib/deepspeed_py_aio_handle.h:14-77, py_ds_aio.cpp:13-45, csrc/aio/common/deepspeed_aio_common.cpp).
// The reference uses libaio io_submit/io_getevents; this build uses a thread pool over pread/pwrite so it has no
// library dependency (libaio is not present on every box) and reaches the same queue-depth parallelism.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#define DSBH_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

struct Chunk {
  int fd;
  char* buf;
  int64_t nbytes;
  int64_t offset;
  bool write;
};

struct AioHandle {
  int64_t block_size;
  int queue_depth;
  bool single_submit, overlap_events;
  int num_threads;
  std::vector<std::thread> workers;
  std::deque<Chunk> queue;
  std::mutex mu;
  std::condition_variable cv, done_cv;
  std::atomic<int64_t> inflight{0};
  std::atomic<int64_t> errors{0};
  std::vector<int> open_fds;
  bool stop = false;
  int64_t completed_ops = 0;

  void worker() {
    for (;;) {
      Chunk c;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || !queue.empty(); });
        if (stop && queue.empty()) return;
        c = queue.front();
        queue.pop_front();
      }
      int64_t done = 0;
      while (done < c.nbytes) {
        ssize_t r = c.write ? pwrite(c.fd, c.buf + done, c.nbytes - done, c.offset + done)
                            : pread(c.fd, c.buf + done, c.nbytes - done, c.offset + done);
        if (r < 0) {
          if (errno == EINTR) continue;
          errors++;
          break;
        }
        if (r == 0) break;  // EOF on read
        done += r;
      }
      if (done < c.nbytes && c.write) errors++;
      if (--inflight == 0) {
        std::lock_guard<std::mutex> lk(mu);
        done_cv.notify_all();
      }
    }
  }
};

int open_file(const char* path, bool write, bool direct) {
  int flags = write ? (O_WRONLY | O_CREAT) : O_RDONLY;
#ifdef O_DIRECT
  if (direct) {
    int fd = open(path, flags | O_DIRECT, 0644);
    if (fd >= 0) return fd;
  }
#endif
  return open(path, flags, 0644);
}

}  // namespace

DSBH_EXPORT void* dsbh_aio_create(int64_t block_size, int queue_depth, int single_submit, int overlap_events, int num_threads) {
  AioHandle* h = new AioHandle();
  h->block_size = block_size > 0 ? block_size : (1 << 20);
  h->queue_depth = queue_depth > 0 ? queue_depth : 8;
  h->single_submit = single_submit != 0;
  h->overlap_events = overlap_events != 0;
  h->num_threads = num_threads > 0 ? num_threads : 1;
  const int nworkers = h->num_threads * (h->overlap_events ? std::min(h->queue_depth, 8) : 1);
  for (int i = 0; i < nworkers; ++i) h->workers.emplace_back([h] { h->worker(); });
  return h;
}
DSBH_EXPORT void dsbh_aio_destroy(void* hp) {
  AioHandle* h = static_cast<AioHandle*>(hp);
  {
    std::lock_guard<std::mutex> lk(h->mu);
    h->stop = true;
  }
  h->cv.notify_all();
  for (auto& t : h->workers) t.join();
  for (int fd : h->open_fds) close(fd);
  delete h;
}
DSBH_EXPORT int64_t dsbh_aio_block_size(void* hp) { return static_cast<AioHandle*>(hp)->block_size; }
DSBH_EXPORT int dsbh_aio_queue_depth(void* hp) { return static_cast<AioHandle*>(hp)->queue_depth; }
DSBH_EXPORT int dsbh_aio_thread_count(void* hp) { return static_cast<AioHandle*>(hp)->num_threads; }

// Returns bytes waited for (>= 0) or -1 if any chunk failed since the last wait.
DSBH_EXPORT int64_t dsbh_aio_wait(void* hp) {
  AioHandle* h = static_cast<AioHandle*>(hp);
  {
    std::unique_lock<std::mutex> lk(h->mu);
    h->done_cv.wait(lk, [&] { return h->inflight.load() == 0; });
  }
  for (int fd : h->open_fds) close(fd);
  h->open_fds.clear();
  int64_t ops = h->completed_ops;
  h->completed_ops = 0;
  if (h->errors.exchange(0) != 0) return -1;
  return ops;
}

static int submit(AioHandle* h, void* buffer, int64_t nbytes, const char* path, bool write, bool async, int64_t file_offset) {
  const bool aligned = (reinterpret_cast<uintptr_t>(buffer) % 4096 == 0) && (nbytes % 4096 == 0) && (h->block_size % 4096 == 0) &&
                       (file_offset % 4096 == 0);
  int fd = open_file(path, write, aligned);
  if (fd < 0) return -1;
  if (!write) {
    struct stat st;
    if (fstat(fd, &st) == 0 && st.st_size < file_offset + nbytes) {
      close(fd);
      return -2;  // file too small
    }
  }
  int64_t nchunks = (nbytes + h->block_size - 1) / h->block_size;
  h->inflight += nchunks;
  {
    std::lock_guard<std::mutex> lk(h->mu);
    for (int64_t i = 0; i < nchunks; ++i) {
      Chunk c;
      c.fd = fd;
      c.buf = static_cast<char*>(buffer) + i * h->block_size;
      c.nbytes = std::min(h->block_size, nbytes - i * h->block_size);
      c.offset = file_offset + i * h->block_size;
      c.write = write;
      h->queue.push_back(c);
    }
    h->open_fds.push_back(fd);
    h->completed_ops += 1;
  }
  h->cv.notify_all();
  if (!async) return dsbh_aio_wait(h) < 0 ? -3 : 0;
  return 0;
}

DSBH_EXPORT int dsbh_aio_pread(void* hp, void* buffer, int64_t nbytes, const char* path, int async, int64_t file_offset) {
  return submit(static_cast<AioHandle*>(hp), buffer, nbytes, path, false, async != 0, file_offset);
}
DSBH_EXPORT int dsbh_aio_pwrite(void* hp, const void* buffer, int64_t nbytes, const char* path, int async, int64_t file_offset) {
  return submit(static_cast<AioHandle*>(hp), const_cast<void*>(buffer), nbytes, path, true, async != 0, file_offset);
}
