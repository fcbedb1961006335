// oracle/shim/mpi.h  --  TEST INFRASTRUCTURE, not product code.
//
// In-process stand-in for the handful of MPI calls on the reference's hot path
// (source/update.cpp:64-104, include/utility.h:284).  No MPI exists in this
// image.  "Ranks" are threads of one process: each thread sets its rank with
// shim_mpi_set_rank() and point-to-point messages go through a mutex-protected
// mailbox keyed by (source, destination, tag).  That is enough to run the
// reference's update::replica_exchange unmodified between replicas.
#pragma once

#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct {
    int dummy;
} MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_SUCCESS 0
#define MPI_LONG_LONG 8
#define MPI_DOUBLE 9
#define MPI_C_BOOL 1
#define MPI_INT 4
#define MPI_SUM 100

namespace shim_mpi {
struct World {
    std::mutex mu;
    std::condition_variable cv;
    std::map<std::tuple<int, int, int>, std::deque<std::vector<char>>> box;
    int size = 1;
};
inline World& world() {
    static World w;
    return w;
}
inline int& my_rank() {
    static thread_local int r = 0;
    return r;
}
inline size_t type_size(MPI_Datatype t) {
    switch (t) {
        case MPI_LONG_LONG: return 8;
        case MPI_DOUBLE: return 8;
        case MPI_INT: return 4;
        default: return 1;
    }
}
inline void put(int src, int dst, int tag, const void* buf, size_t bytes) {
    World& w = world();
    std::lock_guard<std::mutex> lk(w.mu);
    w.box[{src, dst, tag}].emplace_back((const char*)buf, (const char*)buf + bytes);
    w.cv.notify_all();
}
inline void take(int src, int dst, int tag, void* buf, size_t bytes) {
    World& w = world();
    std::unique_lock<std::mutex> lk(w.mu);
    auto key = std::make_tuple(src, dst, tag);
    w.cv.wait(lk, [&] { return !w.box[key].empty(); });
    std::vector<char> msg = std::move(w.box[key].front());
    w.box[key].pop_front();
    std::memcpy(buf, msg.data(), std::min(bytes, msg.size()));
}
}  // namespace shim_mpi

inline void shim_mpi_set_rank(int rank, int size) {
    shim_mpi::my_rank() = rank;
    shim_mpi::world().size = size;
}

inline int MPI_Comm_rank(MPI_Comm, int* rank) {
    *rank = shim_mpi::my_rank();
    return MPI_SUCCESS;
}
inline int MPI_Comm_size(MPI_Comm, int* size) {
    *size = shim_mpi::world().size;
    return MPI_SUCCESS;
}
inline int MPI_Send(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm) {
    shim_mpi::put(shim_mpi::my_rank(), dest, tag, buf, count * shim_mpi::type_size(t));
    return MPI_SUCCESS;
}
inline int MPI_Recv(void* buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm, MPI_Status*) {
    shim_mpi::take(source, shim_mpi::my_rank(), tag, buf, count * shim_mpi::type_size(t));
    return MPI_SUCCESS;
}
inline int MPI_Sendrecv(const void* sbuf, int scount, MPI_Datatype st, int dest, int stag, void* rbuf,
                        int rcount, MPI_Datatype rt, int source, int rtag, MPI_Comm c, MPI_Status* s) {
    MPI_Send(sbuf, scount, st, dest, stag, c);
    return MPI_Recv(rbuf, rcount, rt, source, rtag, c, s);
}
inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
