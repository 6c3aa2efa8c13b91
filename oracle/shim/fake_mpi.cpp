// oracle/shim/fake_mpi.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// Threads-as-ranks implementation of the MPI subset declared in mpi.h.
#include "mpi.h"

#include <pthread.h>
#include <sched.h>

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <exception>
#include <map>
#include <mutex>
#include <stdexcept>
#include <thread>
#include <tuple>
#include <vector>

struct fake_mpi_group {
  std::vector<int> ranks;  // world ranks, in group order
};

struct fake_mpi_request {
  bool isRecv = false;
  bool done = false;
  void* buf = nullptr;
  std::size_t bytes = 0;
  int src = 0, dst = 0, tag = 0;
};

namespace {

struct World {
  int n = 1;
  std::mutex m;
  std::condition_variable cv;
  // barrier
  int arrived = 0;
  unsigned long generation = 0;
  bool aborted = false;
  // collective staging
  std::vector<const void*> ptrs;
  // mailboxes
  std::map<std::tuple<int, int, int>, std::deque<std::vector<char>>> boxes;
};

World g_single;           // used outside fake_mpi_run: one rank
World* g_world = &g_single;
thread_local int tl_rank = 0;

std::size_t typeSize(MPI_Datatype t) {
  switch (t) {
    case MPI_BYTE: return 1;
    case MPI_INT: return 4;
    case MPI_FLOAT: return 4;
    case MPI_DOUBLE: return 8;
    case MPI_LONG_LONG: return 8;
    case MPI_UINT64_T: return 8;
  }
  throw std::runtime_error("fake MPI: unknown datatype");
}

void barrier(World& w) {
  std::unique_lock<std::mutex> lock(w.m);
  if (w.aborted) throw std::runtime_error("fake MPI: aborted");
  const unsigned long gen = w.generation;
  if (++w.arrived == w.n) {
    w.arrived = 0;
    ++w.generation;
    w.cv.notify_all();
    return;
  }
  w.cv.wait(lock, [&] { return w.generation != gen || w.aborted; });
  if (w.generation == gen && w.aborted)
    throw std::runtime_error("fake MPI: aborted");
}

// Publish a pointer per rank, run body() while all are visible, then release.
template <class F>
void exchange(const void* mine, F&& body) {
  World& w = *g_world;
  if (w.ptrs.size() != static_cast<std::size_t>(w.n)) {
    std::lock_guard<std::mutex> lock(w.m);
    if (w.ptrs.size() != static_cast<std::size_t>(w.n)) w.ptrs.resize(w.n);
  }
  w.ptrs[tl_rank] = mine;
  barrier(w);
  body(w.ptrs);
  barrier(w);
}

template <class T>
void reduceTyped(T* acc, const T* src, int count, MPI_Op op) {
  for (int i = 0; i < count; ++i) {
    switch (op) {
      case MPI_MIN: acc[i] = std::min(acc[i], src[i]); break;
      case MPI_MAX: acc[i] = std::max(acc[i], src[i]); break;
      case MPI_SUM: acc[i] = acc[i] + src[i]; break;
      default: throw std::runtime_error("fake MPI: unknown op");
    }
  }
}

void reduceInto(void* acc, const void* src, int count, MPI_Datatype t,
                MPI_Op op) {
  switch (t) {
    case MPI_INT:
      reduceTyped(static_cast<int*>(acc), static_cast<const int*>(src), count, op);
      break;
    case MPI_FLOAT:
      reduceTyped(static_cast<float*>(acc), static_cast<const float*>(src), count, op);
      break;
    case MPI_DOUBLE:
      reduceTyped(static_cast<double*>(acc), static_cast<const double*>(src), count, op);
      break;
    case MPI_LONG_LONG:
      reduceTyped(static_cast<long long*>(acc), static_cast<const long long*>(src), count, op);
      break;
    case MPI_UINT64_T:
      reduceTyped(static_cast<std::uint64_t*>(acc), static_cast<const std::uint64_t*>(src), count, op);
      break;
    default:
      throw std::runtime_error("fake MPI: cannot reduce this datatype");
  }
}

// requires w.m held
bool tryComplete(World& w, fake_mpi_request* r) {
  if (r->done) return true;
  auto it = w.boxes.find(std::make_tuple(r->src, r->dst, r->tag));
  if (it == w.boxes.end() || it->second.empty()) return false;
  std::vector<char>& msg = it->second.front();
  std::memcpy(r->buf, msg.data(), std::min(r->bytes, msg.size()));
  it->second.pop_front();
  r->done = true;
  return true;
}

}  // namespace

extern "C" {

int MPI_Comm_rank(MPI_Comm, int* rank) {
  *rank = tl_rank;
  return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm, int* size) {
  *size = g_world->n;
  return MPI_SUCCESS;
}
int MPI_Comm_group(MPI_Comm, MPI_Group* group) {
  auto* g = new fake_mpi_group;
  g->ranks.resize(g_world->n);
  for (int i = 0; i < g_world->n; ++i) g->ranks[i] = i;
  *group = g;
  return MPI_SUCCESS;
}
int MPI_Group_size(MPI_Group group, int* size) {
  *size = static_cast<int>(group->ranks.size());
  return MPI_SUCCESS;
}
int MPI_Group_rank(MPI_Group group, int* rank) {
  *rank = MPI_UNDEFINED;
  for (std::size_t i = 0; i < group->ranks.size(); ++i)
    if (group->ranks[i] == tl_rank) *rank = static_cast<int>(i);
  return MPI_SUCCESS;
}
int MPI_Group_free(MPI_Group* group) {
  delete *group;
  *group = MPI_GROUP_NULL;
  return MPI_SUCCESS;
}
int MPI_Group_incl(MPI_Group group, int n, const int ranks[], MPI_Group* out) {
  auto* g = new fake_mpi_group;
  for (int i = 0; i < n; ++i) g->ranks.push_back(group->ranks.at(ranks[i]));
  *out = g;
  return MPI_SUCCESS;
}
int MPI_Group_range_incl(MPI_Group group, int n, int ranges[][3],
                         MPI_Group* out) {
  auto* g = new fake_mpi_group;
  for (int r = 0; r < n; ++r) {
    const int first = ranges[r][0], last = ranges[r][1], stride = ranges[r][2];
    if (stride > 0)
      for (int i = first; i <= last; i += stride)
        g->ranks.push_back(group->ranks.at(i));
    else if (stride < 0)
      for (int i = first; i >= last; i += stride)
        g->ranks.push_back(group->ranks.at(i));
  }
  *out = g;
  return MPI_SUCCESS;
}
int MPI_Group_translate_ranks(MPI_Group g1, int n, const int ranks1[],
                              MPI_Group g2, int ranks2[]) {
  for (int i = 0; i < n; ++i) {
    const int world = g1->ranks.at(ranks1[i]);
    ranks2[i] = MPI_UNDEFINED;
    for (std::size_t j = 0; j < g2->ranks.size(); ++j)
      if (g2->ranks[j] == world) ranks2[i] = static_cast<int>(j);
  }
  return MPI_SUCCESS;
}

int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag,
              MPI_Comm, MPI_Request* req) {
  World& w = *g_world;
  const std::size_t bytes = typeSize(type) * static_cast<std::size_t>(count);
  std::vector<char> msg(bytes);
  if (bytes) std::memcpy(msg.data(), buf, bytes);
  {
    std::lock_guard<std::mutex> lock(w.m);
    w.boxes[std::make_tuple(tl_rank, dest, tag)].push_back(std::move(msg));
  }
  w.cv.notify_all();
  auto* r = new fake_mpi_request;
  r->done = true;
  *req = r;
  return MPI_SUCCESS;
}

int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag,
              MPI_Comm, MPI_Request* req) {
  auto* r = new fake_mpi_request;
  r->isRecv = true;
  r->buf = buf;
  r->bytes = typeSize(type) * static_cast<std::size_t>(count);
  r->src = source;
  r->dst = tl_rank;
  r->tag = tag;
  *req = r;
  return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request reqs[], MPI_Status*) {
  World& w = *g_world;
  for (int i = 0; i < count; ++i) {
    if (reqs[i] == MPI_REQUEST_NULL) continue;
    {
      std::unique_lock<std::mutex> lock(w.m);
      w.cv.wait(lock, [&] { return tryComplete(w, reqs[i]) || w.aborted; });
      if (!reqs[i]->done) throw std::runtime_error("fake MPI: aborted");
    }
    delete reqs[i];
    reqs[i] = MPI_REQUEST_NULL;
  }
  return MPI_SUCCESS;
}

int MPI_Waitany(int count, MPI_Request reqs[], int* index, MPI_Status*) {
  World& w = *g_world;
  std::unique_lock<std::mutex> lock(w.m);
  for (;;) {
    bool any = false;
    for (int i = 0; i < count; ++i) {
      if (reqs[i] == MPI_REQUEST_NULL) continue;
      any = true;
      if (tryComplete(w, reqs[i])) {
        delete reqs[i];
        reqs[i] = MPI_REQUEST_NULL;
        *index = i;
        return MPI_SUCCESS;
      }
    }
    if (!any) {
      *index = MPI_UNDEFINED;
      return MPI_SUCCESS;
    }
    if (w.aborted) throw std::runtime_error("fake MPI: aborted");
    w.cv.wait(lock);
  }
}

int MPI_Barrier(MPI_Comm) {
  barrier(*g_world);
  return MPI_SUCCESS;
}

int MPI_Allgather(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
                  int rcount, MPI_Datatype rtype, MPI_Comm) {
  const std::size_t sbytes = typeSize(stype) * scount;
  const std::size_t rbytes = typeSize(rtype) * rcount;
  exchange(sbuf, [&](const std::vector<const void*>& p) {
    for (int r = 0; r < g_world->n; ++r)
      std::memcpy(static_cast<char*>(rbuf) + rbytes * r, p[r],
                  std::min(sbytes, rbytes));
  });
  return MPI_SUCCESS;
}

namespace {
struct VSend {
  const void* buf;
  std::size_t bytes;
};
}  // namespace

int MPI_Allgatherv(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
                   const int rcounts[], const int displs[], MPI_Datatype rtype,
                   MPI_Comm) {
  VSend mine{sbuf, typeSize(stype) * static_cast<std::size_t>(scount)};
  const std::size_t rs = typeSize(rtype);
  exchange(&mine, [&](const std::vector<const void*>& p) {
    for (int r = 0; r < g_world->n; ++r) {
      const VSend* s = static_cast<const VSend*>(p[r]);
      const std::size_t want = rs * static_cast<std::size_t>(rcounts[r]);
      if (want)
        std::memcpy(static_cast<char*>(rbuf) + rs * displs[r], s->buf,
                    std::min(want, s->bytes));
    }
  });
  return MPI_SUCCESS;
}

int MPI_Gather(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
               int rcount, MPI_Datatype rtype, int root, MPI_Comm) {
  const std::size_t sbytes = typeSize(stype) * scount;
  exchange(sbuf, [&](const std::vector<const void*>& p) {
    if (tl_rank != root) return;
    const std::size_t rbytes = typeSize(rtype) * rcount;
    for (int r = 0; r < g_world->n; ++r)
      std::memcpy(static_cast<char*>(rbuf) + rbytes * r, p[r],
                  std::min(sbytes, rbytes));
  });
  return MPI_SUCCESS;
}

int MPI_Gatherv(const void* sbuf, int scount, MPI_Datatype stype, void* rbuf,
                const int rcounts[], const int displs[], MPI_Datatype rtype,
                int root, MPI_Comm) {
  VSend mine{sbuf, typeSize(stype) * static_cast<std::size_t>(scount)};
  exchange(&mine, [&](const std::vector<const void*>& p) {
    if (tl_rank != root) return;
    const std::size_t rs = typeSize(rtype);
    for (int r = 0; r < g_world->n; ++r) {
      const VSend* s = static_cast<const VSend*>(p[r]);
      const std::size_t want = rs * static_cast<std::size_t>(rcounts[r]);
      if (want)
        std::memcpy(static_cast<char*>(rbuf) + rs * displs[r], s->buf,
                    std::min(want, s->bytes));
    }
  });
  return MPI_SUCCESS;
}

int MPI_Allreduce(const void* sbuf, void* rbuf, int count, MPI_Datatype type,
                  MPI_Op op, MPI_Comm) {
  const std::size_t bytes = typeSize(type) * static_cast<std::size_t>(count);
  std::vector<char> acc(bytes);
  exchange(sbuf, [&](const std::vector<const void*>& p) {
    std::memcpy(acc.data(), p[0], bytes);
    for (int r = 1; r < g_world->n; ++r)
      reduceInto(acc.data(), p[r], count, type, op);
  });
  std::memcpy(rbuf, acc.data(), bytes);
  return MPI_SUCCESS;
}

int MPI_Reduce(const void* sbuf, void* rbuf, int count, MPI_Datatype type,
               MPI_Op op, int root, MPI_Comm) {
  const std::size_t bytes = typeSize(type) * static_cast<std::size_t>(count);
  std::vector<char> acc(bytes);
  exchange(sbuf, [&](const std::vector<const void*>& p) {
    if (tl_rank != root) return;
    std::memcpy(acc.data(), p[0], bytes);
    for (int r = 1; r < g_world->n; ++r)
      reduceInto(acc.data(), p[r], count, type, op);
  });
  if (tl_rank == root) std::memcpy(rbuf, acc.data(), bytes);
  return MPI_SUCCESS;
}

int MPI_Abort(MPI_Comm, int code) {
  std::fprintf(stderr, "fake MPI_Abort(%d)\n", code);
  std::abort();
}

}  // extern "C"

void fake_mpi_run(int nranks, const std::function<void(int)>& fn,
                  bool pinCores) {
  if (nranks < 1) throw std::invalid_argument("fake_mpi_run: nranks < 1");
  if (g_world != &g_single)
    throw std::runtime_error("fake_mpi_run: already inside a fake MPI world");
  World world;
  world.n = nranks;
  world.ptrs.resize(nranks);
  g_world = &world;

  std::vector<int> cores;
  if (pinCores) {
    cpu_set_t set;
    CPU_ZERO(&set);
    if (sched_getaffinity(0, sizeof(set), &set) == 0)
      for (int c = 0; c < CPU_SETSIZE; ++c)
        if (CPU_ISSET(c, &set)) cores.push_back(c);
  }

  std::vector<std::exception_ptr> errors(nranks);
  std::vector<std::thread> threads;
  threads.reserve(nranks);
  for (int r = 0; r < nranks; ++r) {
    threads.emplace_back([&, r] {
      tl_rank = r;
      if (!cores.empty()) {
        cpu_set_t one;
        CPU_ZERO(&one);
        CPU_SET(cores[r % cores.size()], &one);
        pthread_setaffinity_np(pthread_self(), sizeof(one), &one);
      }
      try {
        fn(r);
      } catch (...) {
        errors[r] = std::current_exception();
        {
          std::lock_guard<std::mutex> lock(world.m);
          world.aborted = true;
        }
        world.cv.notify_all();
      }
    });
  }
  for (auto& t : threads) t.join();
  g_world = &g_single;
  tl_rank = 0;
  for (auto& e : errors)
    if (e) std::rethrow_exception(e);
}
