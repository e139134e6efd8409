#pragma once
/* mpi::world_rank()/world_size(): rank and size of the process group.
   Parity: reference include/stencil/mpi.hpp:9-27. Backed by stencil::world() (comm.hpp) rather than MPI, so they
   work with no MPI installed (rank 0 / size 1 for a lone process). */
namespace mpi {
int world_rank();
int world_size();
} // namespace mpi
