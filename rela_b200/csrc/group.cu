// group.cu -- rendezvous of the processes that each hold one shard of a striped
// replay (one process per GPU of ONE box, e.g. the ranks of a torchrun job).
//
// No reference counterpart: rela has one host-memory buffer and no
// communication backend. The group is used at SET-UP time only -- to exchange
// CUDA IPC handles of slabs, batch buffers, inboxes and flags, and to agree on
// the schema; the steady-state add / sample / update_priority cycle exchanges
// data through NVLink peer memory and device-side flags and never comes here.
//
// Mechanism: a POSIX shared-memory segment "/rela_b200.<name>" created by rank
// 0 (any stale segment of that name is unlinked first). Joining is a nonce
// hand-shake so that a rank that happens to open a stale segment of an earlier
// job (whose rank 0 is gone and cannot answer) notices and retries by name.
// Collectives are sequence-numbered: all-gather of one fixed-size blob per rank
// through per-rank mailboxes, bracketed by two counter barriers.
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/rela_b200.h"

namespace {

constexpr int kGroupMaxRanks = 16;
constexpr uint64_t kGroupMagic = 0x72656c6162323030ull;  // "relab200"
constexpr size_t kMailbox = 64 * 1024;

struct GroupShm {
  std::atomic<uint64_t> magic;  // set last by rank 0
  std::atomic<uint64_t> session;
  std::atomic<int32_t> world;
  std::atomic<uint64_t> nonce[kGroupMaxRanks];  // written by the joining rank
  std::atomic<uint64_t> ack[kGroupMaxRanks];    // echoed by rank 0
  std::atomic<uint64_t> arrive[kGroupMaxRanks];  // collective sequence numbers
  std::atomic<uint64_t> depart[kGroupMaxRanks];
  uint8_t mail[kGroupMaxRanks][kMailbox];
};

using clk = std::chrono::steady_clock;

inline double since(clk::time_point t0) {
  return std::chrono::duration<double>(clk::now() - t0).count();
}

inline void nap(int spins) {
  if (spins < 200) {
    std::this_thread::yield();
  } else {
    std::this_thread::sleep_for(std::chrono::microseconds(200));
  }
}

int gfail(int code, const char* fmt, const char* a = "", const char* b = "");

}  // namespace

struct rr_group {
  GroupShm* shm = nullptr;
  std::string shm_name;
  int rank = 0, world = 1;
  uint64_t seq = 0;
  double timeout_s = 120.0;
};

// rr_last_error lives in replay.cu; this file reports through the same thread-local text
extern "C" void rr_set_last_error_(const char* text);

namespace {
int gfail(int code, const char* fmt, const char* a, const char* b) {
  char buf[512];
  snprintf(buf, sizeof(buf), fmt, a, b);
  rr_set_last_error_(buf);
  return code;
}

bool wait_all(rr_group* g, std::atomic<uint64_t>* cnt, uint64_t want) {
  const auto t0 = clk::now();
  for (int r = 0; r < g->world; ++r) {
    int spins = 0;
    while (cnt[r].load(std::memory_order_acquire) < want) {
      if (since(t0) > g->timeout_s) return false;
      nap(spins++);
    }
  }
  return true;
}
}  // namespace

extern "C" {

int rr_group_create(rr_group** out, const char* name, int rank, int world) {
  if (out == nullptr || name == nullptr || world < 1 || world > kGroupMaxRanks || rank < 0 || rank >= world)
    return gfail(RR_ERR_ARG, "rr_group_create: bad argument");
  rr_group* g = new rr_group();
  g->rank = rank;
  g->world = world;
  g->shm_name = std::string("/rela_b200.") + name;
  for (char& c : g->shm_name)
    if (c == '/' && &c != &g->shm_name[0]) c = '_';
  if (const char* t = getenv("RELA_B200_GROUP_TIMEOUT")) g->timeout_s = atof(t);
  const auto t0 = clk::now();
  if (rank == 0) {
    shm_unlink(g->shm_name.c_str());  // a stale segment of an earlier job
    const int fd = shm_open(g->shm_name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, sizeof(GroupShm)) != 0) {
      if (fd >= 0) close(fd);
      const std::string nm = g->shm_name;
      delete g;
      return gfail(RR_ERR_STATE, "rr_group_create: cannot create %s: %s", nm.c_str(), strerror(errno));
    }
    void* p = mmap(nullptr, sizeof(GroupShm), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
      delete g;
      return gfail(RR_ERR_STATE, "rr_group_create: mmap failed: %s", strerror(errno));
    }
    g->shm = static_cast<GroupShm*>(p);  // a fresh segment is zero-filled
    g->shm->world.store(world);
    g->shm->session.store((uint64_t)clk::now().time_since_epoch().count() ^ ((uint64_t)getpid() << 32));
    g->shm->magic.store(kGroupMagic, std::memory_order_release);
    // answer the nonces of the joining ranks until all of them are in
    int joined = 1, spins = 0;
    bool in[kGroupMaxRanks] = {true};
    while (joined < world) {
      for (int r = 1; r < world; ++r) {
        const uint64_t n = g->shm->nonce[r].load(std::memory_order_acquire);
        if (n != 0 && g->shm->ack[r].load() != n) {
          g->shm->ack[r].store(n, std::memory_order_release);
          if (!in[r]) {
            in[r] = true;
            ++joined;
          }
        }
      }
      if (since(t0) > g->timeout_s) {
        munmap(g->shm, sizeof(GroupShm));
        shm_unlink(g->shm_name.c_str());
        delete g;
        return gfail(RR_ERR_STATE, "rr_group_create: timed out waiting for the other ranks of group %s", name);
      }
      nap(spins++);
    }
  } else {
    // a nonce no stale segment can already hold an answer for
    const uint64_t nonce = ((uint64_t)clk::now().time_since_epoch().count() * 0x9e3779b97f4a7c15ull) ^
                           ((uint64_t)getpid() << 20) ^ (uint64_t)rank | 1ull;
    int spins = 0;
    for (;;) {
      if (since(t0) > g->timeout_s) {
        delete g;
        return gfail(RR_ERR_STATE, "rr_group_create: timed out joining group %s (is rank 0 running?)", name);
      }
      const int fd = shm_open(g->shm_name.c_str(), O_RDWR, 0600);
      if (fd < 0) {
        nap(spins++);
        continue;
      }
      struct stat st;
      if (fstat(fd, &st) != 0 || (size_t)st.st_size < sizeof(GroupShm)) {  // rank 0 has not sized it yet
        close(fd);
        nap(spins++);
        continue;
      }
      void* p = mmap(nullptr, sizeof(GroupShm), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
      close(fd);
      if (p == MAP_FAILED) {
        nap(spins++);
        continue;
      }
      GroupShm* shm = static_cast<GroupShm*>(p);
      bool ok = false;
      const auto t1 = clk::now();
      int sp2 = 0;
      while (since(t1) < 0.5) {  // a live rank 0 answers within microseconds
        if (shm->magic.load(std::memory_order_acquire) == kGroupMagic && shm->world.load() == world) {
          if (shm->nonce[rank].load() != nonce) shm->nonce[rank].store(nonce, std::memory_order_release);
          if (shm->ack[rank].load(std::memory_order_acquire) == nonce) {
            ok = true;
            break;
          }
        }
        nap(sp2++);
      }
      if (ok) {
        g->shm = shm;
        break;
      }
      munmap(p, sizeof(GroupShm));  // stale or not ready: look the name up again
      nap(spins++);
    }
  }
  *out = g;
  // every rank is in: rank 0 can drop the name (the mapping stays alive)
  int rc = rr_group_barrier(g);
  if (rc != RR_OK) {
    rr_group_destroy(g);
    *out = nullptr;
    return rc;
  }
  if (rank == 0) shm_unlink(g->shm_name.c_str());
  return RR_OK;
}

int rr_group_allgather(rr_group* g, const void* mine, uint64_t bytes, void* all) {
  if (g == nullptr || (bytes > 0 && (mine == nullptr || all == nullptr)) || bytes > kMailbox)
    return gfail(RR_ERR_ARG, "rr_group_allgather: bad argument (at most 64 KB per rank)");
  const uint64_t s = ++g->seq;
  // nobody still reads my mailbox from the previous collective
  if (!wait_all(g, g->shm->depart, s - 1)) return gfail(RR_ERR_STATE, "rr_group_allgather: timed out (a rank left the group?)");
  if (bytes > 0) memcpy(g->shm->mail[g->rank], mine, bytes);
  g->shm->arrive[g->rank].store(s, std::memory_order_release);
  if (!wait_all(g, g->shm->arrive, s)) return gfail(RR_ERR_STATE, "rr_group_allgather: timed out (a rank left the group?)");
  for (int r = 0; r < g->world && bytes > 0; ++r)
    memcpy(static_cast<uint8_t*>(all) + (size_t)r * bytes, g->shm->mail[r], bytes);
  g->shm->depart[g->rank].store(s, std::memory_order_release);
  return RR_OK;
}

int rr_group_barrier(rr_group* g) {
  return rr_group_allgather(g, nullptr, 0, nullptr);
}

int rr_group_rank(const rr_group* g) { return g ? g->rank : -1; }
int rr_group_world(const rr_group* g) { return g ? g->world : 0; }

int rr_group_destroy(rr_group* g) {
  if (g == nullptr) return RR_OK;
  if (g->shm) munmap(g->shm, sizeof(GroupShm));
  delete g;
  return RR_OK;
}

}  // extern "C"
