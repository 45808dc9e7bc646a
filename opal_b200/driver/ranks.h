// ranks.h -- one process per GPU on one node, the stand-in for `mpirun -n N` + rsmpi (main.rs:49-51).
// The data path between slabs (halo, migration) is NCCL inside libopal_b200.so; the host side only needs the small
// collectives of the output code (min/max/sum all-reduce, gather to rank 0: hgram.rs:199-204, yee.rs:755-760,
// mod.rs:232-239) and a way to hand the NCCL unique id to every rank.  Both go through an anonymous shared mapping
// created before fork(): per-rank staging slots plus a process-shared barrier.
#pragma once
#include <functional>
#include <vector>

#include "output.h"

namespace opal_driver {

class Ranks : public Reducer {
public:
    // joins this rank's context to the ring communicator (opal_comm_unique_id on rank 0, opal_comm_init everywhere)
    virtual void connect(opal_ctx* ctx) = 0;
    // True when the library has no communicator of its own (opal_comm_unique_id answers OPAL_ENCCL: the CPU oracle build, or a
    // box without NCCL).  The driver then moves leavers and halo strips itself, call by call, as a Rust driver that keeps MPI
    // for its p2p would (opal_migrants_take/give = mod.rs:296-373, opal_halo_take/give = yee.rs:353-409).
    virtual bool host_staged() const { return false; }
    // Ring exchange of two byte strings with the neighbours (rank -+ 1) mod n: what the left neighbour passed as to_right
    // arrives in from_left, what the right neighbour passed as to_left arrives in from_right.
    virtual void ring_exchange(const std::vector<char>& to_left, const std::vector<char>& to_right, std::vector<char>& from_left,
                               std::vector<char>& from_right) {
        from_left = to_right;   // one rank: its own neighbour on both sides
        from_right = to_left;
    }
};

// runs body(ranks) in n processes (n == 1: in this process) and returns the largest exit code
int launch_ranks(int n, const std::function<int(Ranks&)>& body);

}  // namespace opal_driver
