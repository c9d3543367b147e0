// mpi.h -- single-rank stand-in so that the reference's unmodified driver (PIC_2D.cpp, which includes <mpi.h>) builds on
// a box without MPI.  World size is 1, rank 0; a nonblocking send is delivered to the receive posted with the same tag
// when MPI_Waitall runs (the reference posts all receives, then all sends, then waits: PIC_2D.cpp:200-248), which is how
// same-rank neighbour tiles exchange guard strips.  Only what that driver uses is provided.  Header-only.
#ifndef MINIPIC_SHIM_MPI_H
#define MINIPIC_SHIM_MPI_H
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
struct MPI_Status { int MPI_SOURCE, MPI_TAG, MPI_ERROR; };
#define MPI_COMM_WORLD 0
#define MPI_BYTE 1
#define MPI_INT 4
#define MPI_MAX 0
#define MPI_IN_PLACE ((void*)1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_SUCCESS 0

namespace minipic_shim {
struct Posted { void* buf; size_t bytes; };
inline std::map<int, Posted>& recvs() { static std::map<int, Posted> m; return m; }
inline std::map<int, std::vector<unsigned char>>& sends() { static std::map<int, std::vector<unsigned char>> m; return m; }
inline size_t width(MPI_Datatype t) { return t == MPI_INT ? sizeof(int) : 1; }
}  // namespace minipic_shim

inline int MPI_Init(int*, char***) { return MPI_SUCCESS; }
inline int MPI_Finalize() { return MPI_SUCCESS; }
inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return MPI_SUCCESS; }
inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
inline int MPI_Allreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm) { return MPI_SUCCESS; }   // one rank: identity
inline int MPI_Irecv(void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm, MPI_Request* rq) {
    minipic_shim::recvs()[tag] = { buf, (size_t)count * minipic_shim::width(t) };
    if (rq) *rq = tag;
    return MPI_SUCCESS;
}
inline int MPI_Isend(const void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm, MPI_Request* rq) {
    const unsigned char* p = (const unsigned char*)buf;
    minipic_shim::sends()[tag].assign(p, p + (size_t)count * minipic_shim::width(t));
    if (rq) *rq = tag;
    return MPI_SUCCESS;
}
inline int MPI_Waitall(int, MPI_Request*, MPI_Status*) {
    auto& rs = minipic_shim::recvs();
    auto& ss = minipic_shim::sends();
    for (auto& kv : ss) {
        auto it = rs.find(kv.first);
        if (it == rs.end() || it->second.bytes != kv.second.size()) {
            std::fprintf(stderr, "mpi shim: send with tag %d has no matching receive\n", kv.first);
            std::abort();
        }
        std::memcpy(it->second.buf, kv.second.data(), kv.second.size());
        rs.erase(it);
    }
    if (!rs.empty()) { std::fprintf(stderr, "mpi shim: %zu receives never matched\n", rs.size()); std::abort(); }
    ss.clear();
    return MPI_SUCCESS;
}
inline int MPI_Send(const void*, int, MPI_Datatype, int, int, MPI_Comm) { std::fprintf(stderr, "mpi shim: MPI_Send at world size 1\n"); std::abort(); }
inline int MPI_Recv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*) { std::fprintf(stderr, "mpi shim: MPI_Recv at world size 1\n"); std::abort(); }
#endif
