/* Single-rank MPI stand-in, used ONLY to build the unmodified reference (mengshanfeng/stencil-2) on machines with no
   MPI installation (this image has none). It is not part of the product library. Scope: exactly the calls the
   reference makes. One process is rank 0 of a world of size 1; collectives are copies; point-to-point
   is a tag-keyed FIFO of self-messages, enough for 1 rank x N GPUs — the reference's cudaMemcpyPeer (--peer),
   same-GPU kernel (--kernel) and staged self-send (--staged) paths. The colocated (CUDA IPC) and CUDA-aware paths
   need two real ranks and cannot run under this shim. */
#ifndef STENCIL_BASELINE_MPI_SHIM_H
#define STENCIL_BASELINE_MPI_SHIM_H

#include <climits>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <time.h>
#include <unistd.h>
#include <vector>

struct mpi_shim_comm {
  int id;
};
typedef mpi_shim_comm *MPI_Comm; /* pointer-typed like Open MPI: the reference assigns nullptr to one */
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Info;
struct MPI_Status {
  int MPI_SOURCE, MPI_TAG, MPI_ERROR;
  int shim_bytes; /* filled by MPI_Iprobe for MPI_Get_count */
};

enum { MPI_SUCCESS = 0 };
enum { MPI_BYTE = 1, MPI_CHAR = 2, MPI_INT = 3, MPI_FLOAT = 4, MPI_DOUBLE = 5, MPI_UINT64_T = 6, MPI_INT64_T = 7, MPI_UNSIGNED = 8 };
enum { MPI_MAX = 1, MPI_MIN = 2, MPI_SUM = 3 };
enum { MPI_COMM_TYPE_SHARED = 1 };
enum { MPI_TAG_UB = 1 };
enum { MPI_THREAD_SINGLE = 0, MPI_THREAD_FUNNELED = 1, MPI_THREAD_SERIALIZED = 2, MPI_THREAD_MULTIPLE = 3 };
#define MPI_INFO_NULL 0
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_IN_PLACE ((void *)-1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_COMM_NULL ((MPI_Comm)0)

namespace mpi_shim {
inline mpi_shim_comm &world_storage() {
  static mpi_shim_comm w{0};
  return w;
}
inline size_t size_of(MPI_Datatype t) {
  switch (t) {
  case MPI_BYTE:
  case MPI_CHAR: return 1;
  case MPI_INT:
  case MPI_FLOAT:
  case MPI_UNSIGNED: return 4;
  default: return 8;
  }
}
struct Pending {
  bool isRecv = false, done = false;
  void *buf = nullptr;
  size_t bytes = 0;
  int tag = 0;
};
inline std::map<int, std::deque<std::vector<unsigned char>>> &mailbox() {
  static std::map<int, std::deque<std::vector<unsigned char>>> m;
  return m;
}
inline bool try_complete(Pending *p) {
  if (p->done) return true;
  auto it = mailbox().find(p->tag);
  if (it == mailbox().end() || it->second.empty()) return false;
  std::vector<unsigned char> &m = it->second.front();
  std::memcpy(p->buf, m.data(), m.size() < p->bytes ? m.size() : p->bytes);
  it->second.pop_front();
  p->done = true;
  return true;
}
} // namespace mpi_shim

typedef mpi_shim::Pending *MPI_Request;
#define MPI_REQUEST_NULL ((MPI_Request)0)
#define MPI_COMM_WORLD (&mpi_shim::world_storage())

inline int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
inline int MPI_Init_thread(int *, char ***, int required, int *provided) {
  if (provided) *provided = required;
  return MPI_SUCCESS;
}
inline int MPI_Finalize() { return MPI_SUCCESS; }
inline int MPI_Initialized(int *flag) {
  *flag = 1; /* nothing to initialise */
  return MPI_SUCCESS;
}
inline int MPI_Finalized(int *flag) {
  *flag = 0;
  return MPI_SUCCESS;
}
inline int MPI_Comm_dup(MPI_Comm c, MPI_Comm *out) {
  *out = c;
  return MPI_SUCCESS;
}
inline int MPI_Comm_rank(MPI_Comm, int *r) {
  *r = 0;
  return MPI_SUCCESS;
}
inline int MPI_Comm_size(MPI_Comm, int *s) {
  *s = 1;
  return MPI_SUCCESS;
}
inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
inline double MPI_Wtime() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return double(ts.tv_sec) + 1e-9 * double(ts.tv_nsec);
}
inline int MPI_Comm_split_type(MPI_Comm c, int, int, MPI_Info, MPI_Comm *out) {
  *out = c;
  return MPI_SUCCESS;
}
inline int MPI_Comm_free(MPI_Comm *) { return MPI_SUCCESS; }
inline int MPI_Comm_get_attr(MPI_Comm, int, void *attr, int *flag) {
  static int tagUb = INT_MAX;
  *static_cast<int **>(attr) = &tagUb;
  *flag = 1;
  return MPI_SUCCESS;
}
inline int MPI_Get_processor_name(char *name, int *len) {
  if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1) != 0) std::strcpy(name, "localhost");
  *len = int(std::strlen(name));
  return MPI_SUCCESS;
}
inline int mpi_shim_copy(const void *s, void *r, int count, MPI_Datatype t) {
  if (s != MPI_IN_PLACE && s != r) std::memcpy(r, s, size_t(count) * mpi_shim::size_of(t));
  return MPI_SUCCESS;
}
inline int MPI_Reduce(const void *s, void *r, int count, MPI_Datatype t, MPI_Op, int, MPI_Comm) { return mpi_shim_copy(s, r, count, t); }
inline int MPI_Allreduce(const void *s, void *r, int count, MPI_Datatype t, MPI_Op, MPI_Comm) { return mpi_shim_copy(s, r, count, t); }
inline int MPI_Allgather(const void *s, int sc, MPI_Datatype st, void *r, int, MPI_Datatype, MPI_Comm) { return mpi_shim_copy(s, r, sc, st); }
inline int MPI_Gather(const void *s, int sc, MPI_Datatype st, void *r, int, MPI_Datatype, int, MPI_Comm) { return mpi_shim_copy(s, r, sc, st); }
inline int MPI_Allgatherv(const void *s, int sc, MPI_Datatype st, void *r, const int *, const int *displs, MPI_Datatype rt, MPI_Comm) {
  return mpi_shim_copy(s, static_cast<char *>(r) + size_t(displs ? displs[0] : 0) * mpi_shim::size_of(rt), sc, st);
}
inline int MPI_Bcast(void *, int, MPI_Datatype, int, MPI_Comm) { return MPI_SUCCESS; }

inline int MPI_Isend(const void *buf, int count, MPI_Datatype t, int, int tag, MPI_Comm, MPI_Request *req) {
  const unsigned char *p = static_cast<const unsigned char *>(buf);
  mpi_shim::mailbox()[tag].emplace_back(p, p + size_t(count) * mpi_shim::size_of(t));
  mpi_shim::Pending *q = new mpi_shim::Pending;
  q->done = true;
  *req = q;
  return MPI_SUCCESS;
}
inline int MPI_Irecv(void *buf, int count, MPI_Datatype t, int, int tag, MPI_Comm, MPI_Request *req) {
  mpi_shim::Pending *q = new mpi_shim::Pending;
  q->isRecv = true;
  q->buf = buf;
  q->bytes = size_t(count) * mpi_shim::size_of(t);
  q->tag = tag;
  *req = q;
  return MPI_SUCCESS;
}
inline int MPI_Test(MPI_Request *req, int *flag, MPI_Status *) {
  if (!*req) {
    *flag = 1;
    return MPI_SUCCESS;
  }
  *flag = mpi_shim::try_complete(*req) ? 1 : 0;
  if (*flag) {
    delete *req;
    *req = MPI_REQUEST_NULL;
  }
  return MPI_SUCCESS;
}
inline int MPI_Wait(MPI_Request *req, MPI_Status *st) {
  int flag = 0;
  while (!flag) MPI_Test(req, &flag, st); /* a lone rank can only wait for something it already sent */
  return MPI_SUCCESS;
}
inline int MPI_Iprobe(int, int tag, MPI_Comm, int *flag, MPI_Status *st) {
  auto it = mpi_shim::mailbox().find(tag);
  *flag = (it != mpi_shim::mailbox().end() && !it->second.empty()) ? 1 : 0;
  if (*flag && st) {
    st->MPI_SOURCE = 0;
    st->MPI_TAG = tag;
    st->MPI_ERROR = MPI_SUCCESS;
    st->shim_bytes = int(it->second.front().size());
  }
  return MPI_SUCCESS;
}
inline int MPI_Get_count(const MPI_Status *st, MPI_Datatype t, int *count) {
  *count = int(size_t(st->shim_bytes) / mpi_shim::size_of(t));
  return MPI_SUCCESS;
}
inline int MPI_Send(const void *buf, int count, MPI_Datatype t, int dst, int tag, MPI_Comm c) {
  MPI_Request r;
  MPI_Isend(buf, count, t, dst, tag, c, &r);
  return MPI_Wait(&r, MPI_STATUS_IGNORE);
}
inline int MPI_Recv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *st) {
  MPI_Request r;
  MPI_Irecv(buf, count, t, src, tag, c, &r);
  return MPI_Wait(&r, st);
}

#endif
