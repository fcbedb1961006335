// dqmcb/update.h -- host mirror of `namespace update` (reference include/update.h, source/update.cpp):
// local_update, partner_rank, replica_exchange with the reference's signatures.
//
// local_update is the teacher-forced single-slice entry: it takes the caller's Gtt[l+1], runs the
// batched Metropolis kernel for slice l with draws taken from `rng` exactly as the reference takes
// them, and hands G / fields / acceptance back.  Whole sweeps should go through DQMC::sweep_* (no
// per-slice host round trip).
//
// replica_exchange keeps the reference's protocol (fields both ways, cross actions, lower rank
// decides, verdict to the partner, rebuild on reject).  The reference hard-wires MPI_COMM_WORLD; here
// the four message kinds go through update::Transport so the same code runs over in-process ranks
// (tests), MPI (-DDQMCB_USE_MPI) or any user transport.  The multi-GPU production path
// (dqmc_b200/host.py, NCCL over NVLink) implements the same protocol on int8 device buffers.
#pragma once
#include <algorithm>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "dqmc.h"
#include "field.h"
#include "model.h"
#include "stackngf.h"
#include "utility.h"
#ifdef DQMCB_USE_MPI
#include <mpi.h>
#endif

namespace update {

// ---- transport ------------------------------------------------------------------------------------
struct Transport {
    virtual ~Transport() = default;
    virtual void send(int self, int dst, int tag, const void* buf, size_t bytes) = 0;
    virtual void recv(int self, int src, int tag, void* buf, size_t bytes) = 0;
    virtual void sendrecv(int self, int partner, int tag, const void* sbuf, void* rbuf, size_t bytes) {
        send(self, partner, tag, sbuf, bytes);
        recv(self, partner, tag, rbuf, bytes);
    }
};

// ranks = threads of one process; buffered sends, blocking receives
class LocalTransport : public Transport {
    std::mutex m_;
    std::condition_variable cv_;
    std::map<std::tuple<int, int, int>, std::vector<std::vector<unsigned char>>> box_;  // (src, dst, tag) -> queue

   public:
    void send(int self, int dst, int tag, const void* buf, size_t bytes) override {
        std::lock_guard<std::mutex> lk(m_);
        const unsigned char* b = static_cast<const unsigned char*>(buf);
        box_[{self, dst, tag}].emplace_back(b, b + bytes);
        cv_.notify_all();
    }
    void recv(int self, int src, int tag, void* buf, size_t bytes) override {
        std::unique_lock<std::mutex> lk(m_);
        auto key = std::make_tuple(src, self, tag);
        cv_.wait(lk, [&] { return !box_[key].empty(); });
        auto& q = box_[key];
        std::memcpy(buf, q.front().data(), std::min(bytes, q.front().size()));
        q.erase(q.begin());
    }
};

#ifdef DQMCB_USE_MPI
class MpiTransport : public Transport {
   public:
    void send(int, int dst, int tag, const void* buf, size_t bytes) override {
        MPI_Send(buf, (int)bytes, MPI_BYTE, dst, tag, MPI_COMM_WORLD);
    }
    void recv(int, int src, int tag, void* buf, size_t bytes) override {
        MPI_Recv(buf, (int)bytes, MPI_BYTE, src, tag, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    }
    void sendrecv(int, int partner, int tag, const void* sbuf, void* rbuf, size_t bytes) override {
        MPI_Sendrecv(sbuf, (int)bytes, MPI_BYTE, partner, tag, rbuf, (int)bytes, MPI_BYTE, partner, tag, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    }
};
#endif

inline Transport*& transport() {
    static Transport* t = nullptr;
    return t;
}

// ---- source/update.cpp:5-32 ---------------------------------------------------------------------------
inline void local_update(utility::random& rng, AttractiveHubbard& model, std::vector<GF>& greens, int l, double& acc_rate) {
    dqmcb_ctx* ctx = model.ctx();
    const int W = model.n_walkers(), nv = model.ns();
    model.sync_fields_to_device();
    model.upload_G(greens[0].Gtt[l + 1]);
    std::vector<uint8_t> perm, k;
    std::vector<double> u;
    dqmcb::draw_tape(rng.get_generator(), nv, 1, perm, k, u);
    for (int w = 1; w < W; ++w) {  // batched walkers beyond the reference's single chain replay walker 0's draws
        perm.insert(perm.end(), perm.begin(), perm.begin() + nv);
        k.insert(k.end(), k.begin(), k.begin() + nv);
        u.insert(u.end(), u.begin(), u.begin() + nv);
    }
    std::vector<double> a0(W), a1(W);
    dqmcb::check(dqmcb_get_stats(ctx, a0.data(), nullptr, nullptr), ctx, "get_stats");
    dqmcb::check(dqmcb_set_tape(ctx, 1, perm.data(), k.data(), u.data()), ctx, "set_tape");
    dqmcb::check(dqmcb_local_update(ctx, l), ctx, "local_update");
    dqmcb::check(dqmcb_get_stats(ctx, a1.data(), nullptr, nullptr), ctx, "get_stats");
    model.download_G(greens[0].Gtt[l + 1]);
    model.refresh_fields_from_device();
    // get_stats reports accepted / (N nt); one slice's accepted / N is the reference's acc_rate (update.cpp:31)
    acc_rate = (a1[0] - a0[0]) * model.nt();
}

// ---- source/update.cpp:34-45 ---------------------------------------------------------------------------
inline int partner_rank(const int rank, const int world_size, const int exchange_attempt) {
    return dqmcb_partner_rank(rank, world_size, exchange_attempt);
}

// ---- source/update.cpp:47-117 --------------------------------------------------------------------------
inline void replica_exchange(int rank, int world_size, utility::random& rng, int& exchange_attempt, int& exchange_accepted,
                             AttractiveHubbard& model, DQMC& sim, std::vector<GF>& greens, std::vector<LDRStack>& propagation_stacks) {
    Transport* tp = transport();
    if (!tp) throw std::logic_error("update::replica_exchange: no transport set (update::transport() = &my_transport)");
    ++exchange_attempt;
    const int partner = partner_rank(rank, world_size, exchange_attempt);
    if (partner < 0 || partner >= world_size) return;

    // swap field configurations with the partner (int64 nt x N, column-major: arma::imat)
    const arma::imat mine = model.fields().fields();
    arma::imat theirs(mine.n_rows, mine.n_cols);
    tp->sendrecv(rank, partner, 0, mine.memptr(), theirs.memptr(), mine.n_elem * sizeof(long long));

    const double S_own = model.global_action(greens);  // S_r({s}_r)
    auto rebuild = [&](const arma::imat& f) {
        model.fields().set_fields(f);
        for (int flv = 0; flv < model.n_flavor(); ++flv) {
            propagation_stacks[flv] = sim.init_stacks(flv);
            greens[flv] = sim.init_greenfunctions(propagation_stacks[flv]);
        }
    };
    rebuild(theirs);
    const double S_cross = model.global_action(greens);  // S_r({s}_partner)

    double S_cross_partner = 0.0, S_own_partner = 0.0;
    tp->sendrecv(rank, partner, 1, &S_cross, &S_cross_partner, sizeof(double));
    tp->sendrecv(rank, partner, 2, &S_own, &S_own_partner, sizeof(double));

    bool accept = false;
    if (rank < partner) {  // the lower rank decides
        const double dS = (S_cross + S_cross_partner) - (S_own + S_own_partner);
        accept = rng.bernoulli(std::min(1.0, std::exp(-dS)));
        if (rank == 0) exchange_accepted += accept;
        const unsigned char v = accept ? 1 : 0;
        tp->send(rank, partner, 3, &v, 1);
    } else {
        unsigned char v = 0;
        tp->recv(rank, partner, 3, &v, 1);
        accept = v != 0;
    }
    if (!accept) rebuild(mine);  // the trial overwrote the state: restore
}

}  // namespace update
