// ranks.cpp -- see ranks.h
#include "ranks.h"

#include <pthread.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace opal_driver {

namespace {

class SoloRanks : public Ranks {
public:
    void connect(opal_ctx*) override {}
};

constexpr size_t CHUNK = 1u << 20;      // doubles per rank and exchange round (8 MiB)

struct Shared {
    pthread_barrier_t barrier;
    unsigned char nccl_id[128];
    int id_status;
    size_t sizes[64];
};

class ShmRanks : public Ranks {
public:
    ShmRanks(int rank, int n, Shared* sh, double* slots) : rank_(rank), n_(n), sh_(sh), slots_(slots) {}
    int rank() const override { return rank_; }
    int nranks() const override { return n_; }

    void connect(opal_ctx* ctx) override {
        if (rank_ == 0) sh_->id_status = (int)OPAL_FN(comm_unique_id)(sh_->nccl_id);
        wait();
        if (sh_->id_status == OPAL_ENCCL) {      // no communicator in the library: host-staged exchange (ranks.h)
            staged_ = true;
            wait();
            return;
        }
        if (sh_->id_status != OPAL_OK) throw std::runtime_error("opal_comm_unique_id failed");
        unsigned char id[128];
        std::memcpy(id, sh_->nccl_id, sizeof id);
        const opal_status st = OPAL_FN(comm_init)(ctx, id);
        if (st != OPAL_OK) throw std::runtime_error(std::string("opal_comm_init failed: ") + OPAL_FN(last_error)(ctx));
        wait();
    }

    bool host_staged() const override { return staged_; }
    void ring_exchange(const std::vector<char>& to_left, const std::vector<char>& to_right, std::vector<char>& from_left,
                       std::vector<char>& from_right) override {
        const uint64_t hdr[2] = {to_left.size(), to_right.size()};
        if (sizeof hdr + hdr[0] + hdr[1] > CHUNK * sizeof(double)) throw std::runtime_error("ring_exchange: message larger than the staging slot");
        char* mine = reinterpret_cast<char*>(slot(rank_));
        std::memcpy(mine, hdr, sizeof hdr);
        if (hdr[0]) std::memcpy(mine + sizeof hdr, to_left.data(), hdr[0]);
        if (hdr[1]) std::memcpy(mine + sizeof hdr + hdr[0], to_right.data(), hdr[1]);
        wait();
        uint64_t h[2];
        const char* l = reinterpret_cast<const char*>(slot((rank_ - 1 + n_) % n_));       // yee.rs:309
        std::memcpy(h, l, sizeof h);
        from_left.assign(l + sizeof h + h[0], l + sizeof h + h[0] + h[1]);
        const char* r = reinterpret_cast<const char*>(slot((rank_ + 1) % n_));            // yee.rs:310
        std::memcpy(h, r, sizeof h);
        from_right.assign(r + sizeof h, r + sizeof h + h[0]);
        wait();
    }

    void min(double* v, int n) override { reduce(v, (size_t)n, 0); }
    void max(double* v, int n) override { reduce(v, (size_t)n, 1); }
    void sum(double* v, size_t n) override { reduce(v, n, 2); }
    void sum_i64(int64_t* v, int n) override {
        // exact in f64 below 2^53
        std::vector<double> d(v, v + n);
        reduce(d.data(), (size_t)n, 2);
        for (int i = 0; i < n; i++) v[i] = (int64_t)d[(size_t)i];
    }

    std::vector<double> gather(const std::vector<double>& mine) override {
        sh_->sizes[rank_] = mine.size();
        wait();
        std::vector<size_t> sizes(sh_->sizes, sh_->sizes + n_);
        wait();
        std::vector<double> all;
        for (int src = 0; src < n_; src++) {
            for (size_t off = 0; off < sizes[(size_t)src]; off += CHUNK) {
                const size_t m = std::min(CHUNK, sizes[(size_t)src] - off);
                if (rank_ == src) std::memcpy(slot(src), mine.data() + off, m * sizeof(double));
                wait();
                if (rank_ == 0) all.insert(all.end(), slot(src), slot(src) + m);
                wait();
            }
        }
        return all;
    }

private:
    double* slot(int r) { return slots_ + (size_t)r * CHUNK; }
    void wait() { pthread_barrier_wait(&sh_->barrier); }
    void reduce(double* v, size_t n, int op) {
        for (size_t off = 0; off < n; off += CHUNK) {
            const size_t m = std::min(CHUNK, n - off);
            std::memcpy(slot(rank_), v + off, m * sizeof(double));
            wait();
            for (size_t i = 0; i < m; i++) {
                double acc = slot(0)[i];                    // rank order: every rank computes the same bits
                for (int r = 1; r < n_; r++) {
                    const double x = slot(r)[i];
                    acc = op == 0 ? std::min(acc, x) : (op == 1 ? std::max(acc, x) : acc + x);
                }
                v[off + i] = acc;
            }
            wait();
        }
    }
    int rank_, n_;
    bool staged_ = false;
    Shared* sh_;
    double* slots_;
};

}  // namespace

int launch_ranks(int n, const std::function<int(Ranks&)>& body) {
    if (n <= 1) {
        SoloRanks solo;
        return body(solo);
    }
    if (n > 64) throw std::runtime_error("at most 64 ranks per node");
    const size_t bytes = sizeof(Shared) + 64 + (size_t)n * CHUNK * sizeof(double);
    void* mem = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (mem == MAP_FAILED) { std::perror("mmap"); return 1; }
    Shared* sh = static_cast<Shared*>(mem);
    double* slots = reinterpret_cast<double*>(static_cast<char*>(mem) + (sizeof(Shared) + 63) / 64 * 64);
    pthread_barrierattr_t attr;
    pthread_barrierattr_init(&attr);
    pthread_barrierattr_setpshared(&attr, PTHREAD_PROCESS_SHARED);
    pthread_barrier_init(&sh->barrier, &attr, (unsigned)n);
    std::fflush(stdout);
    std::fflush(stderr);
    std::vector<pid_t> pids;
    for (int r = 0; r < n; r++) {                     // fork BEFORE any CUDA call: each child creates its own context
        const pid_t pid = fork();
        if (pid < 0) { std::perror("fork"); break; }
        if (pid == 0) {
            ShmRanks ranks(r, n, sh, slots);
            int rc = 1;
            try { rc = body(ranks); } catch (...) { rc = 1; }
            std::fflush(stdout);
            std::fflush(stderr);
            _exit(rc);
        }
        pids.push_back(pid);
    }
    int worst = (int)pids.size() == n ? 0 : 1;
    size_t left = pids.size();
    while (left > 0) {
        int status = 0;
        const pid_t done = wait(&status);
        if (done < 0) break;
        left--;
        const int rc = WIFEXITED(status) ? WEXITSTATUS(status) : 128 + (WIFSIGNALED(status) ? WTERMSIG(status) : 0);
        if (rc != 0) {
            worst = std::max(worst, rc);
            for (pid_t p : pids) if (p != done) kill(p, SIGTERM);   // a failed rank would leave the others at a barrier
        }
    }
    munmap(mem, bytes);
    return worst;
}

}  // namespace opal_driver
