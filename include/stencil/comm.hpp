#pragma once
/* stencil::Comm — the process-group "plumbing" this library needs from a launcher.

   The reference hard-wires MPI (mpi.hpp, mpi_topology.hpp, every MPI_* call listed in SURVEY 2.4). On a
   single NVSwitch box the data path here never touches the host, so all that is left for a process group to do
   is set-up traffic (hostnames, CUDA ids, IPC handles, placement broadcast), statistics reductions, and the
   cross-node *staged* fallback. That is a small interface, so it is abstract:

     SelfComm   one process (possibly driving several GPUs)          — default
     TcpComm    N processes, star over TCP through rank 0            — env: RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT
                (what `torchrun --no-python ./jacobi3d ...` or bench.py under torchrun gives us)
     user       anything else (e.g. a torch.distributed- or MPI-backed subclass) via set_world()

   Point-to-point messages are tagged byte blobs with eager buffering (send never blocks on the receiver), which is
   exactly the Isend/Irecv/Test subset the reference's transports use (tx_cuda.cuh:239-281,599-768). */
#include <cstddef>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace stencil {

class Comm {
public:
  virtual ~Comm() {}
  virtual int rank() const = 0;
  virtual int size() const = 0;
  virtual const char *name() const = 0;
  /* true if send()/try_recv() accept DEVICE pointers (a CUDA-aware MPI underneath): enables MethodFlags::CudaAwareMpi */
  virtual bool device_buffers() const { return false; }

  /* eager tagged send; returns once the payload is buffered or on the wire */
  virtual void send(int dst, int tag, const void *buf, size_t bytes) = 0;
  /* Block until every buffer handed to send() so far may be reused. Communicators that copy the payload inside
     send() (all the built-in ones for host pointers) have nothing to do. A communicator that sends IN PLACE — MpiComm
     with device pointers — completes its outstanding sends here; callers invoke it only after they have posted or
     finished their own receives, so two ranks sending each other rendezvous-sized messages cannot block each other. */
  virtual void flush_sends() {}
  /* non-blocking: true and fills buf iff a matching (src, tag) message of exactly `bytes` has arrived */
  virtual bool try_recv(int src, int tag, void *buf, size_t bytes) = 0;
  /* blocking receive (spins on try_recv with a timeout; fatal on timeout) */
  virtual void recv(int src, int tag, void *buf, size_t bytes);

  /* collectives: default implementations are built on send/recv through rank 0 */
  virtual void barrier();
  virtual void allgather(const void *sendbuf, void *recvbuf, size_t bytesPerRank);
  virtual void bcast(void *buf, size_t bytes, int root);
  /* variable-size allgather; counts[r] = bytes contributed by rank r (known to all) */
  virtual void allgatherv(const void *sendbuf, void *recvbuf, const std::vector<size_t> &counts);

  double allreduce_max(double v);
  uint64_t allreduce_sum(uint64_t v);
  std::string hostname() const;
  /* wall clock in seconds (MPI_Wtime stand-in) */
  static double wtime();
};

/* single-process communicator; self-sends go through an in-memory mailbox */
std::unique_ptr<Comm> make_self_comm();
/* TCP star; rank 0 listens on addr:port. Blocks until all `size` ranks joined. */
std::unique_ptr<Comm> make_tcp_comm(int rank, int size, const std::string &addr, int port);

/* MPI-backed communicator (src/mpi_comm.cpp); only in builds configured with -DUSE_MPI=ON, otherwise returns null.
   Initialises MPI if nobody has; device_buffers() follows -DUSE_CUDA_AWARE_MPI=ON. */
std::unique_ptr<Comm> make_mpi_comm();

/* process-wide communicator. First use initialises from the environment:
   WORLD_SIZE>1 with RANK and MASTER_ADDR set -> TcpComm on STENCIL_PORT (default MASTER_PORT+17); else SelfComm. */
Comm &world();
void set_world(std::unique_ptr<Comm> c);
/* tear down the global communicator (MPI_Finalize stand-in); safe to call twice */
void finalize_world();

} // namespace stencil
