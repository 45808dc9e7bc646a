// ranks.h -- one process per GPU on one node, the stand-in for `mpirun -n N` + rsmpi (main.rs:49-51).
// The data path between slabs (halo, migration) is NCCL inside libopal_b200.so; the host side only needs the small
// collectives of the output code (min/max/sum all-reduce, gather to rank 0: hgram.rs:199-204, yee.rs:755-760,
// mod.rs:232-239) and a way to hand the NCCL unique id to every rank.  Both go through an anonymous shared mapping
// created before fork(): per-rank staging slots plus a process-shared barrier.
#pragma once
#include <functional>

#include "output.h"

namespace opal_driver {

class Ranks : public Reducer {
public:
    // joins this rank's context to the ring communicator (opal_comm_unique_id on rank 0, opal_comm_init everywhere)
    virtual void connect(opal_ctx* ctx) = 0;
};

// runs body(ranks) in n processes (n == 1: in this process) and returns the largest exit code
int launch_ranks(int n, const std::function<int(Ranks&)>& body);

}  // namespace opal_driver
