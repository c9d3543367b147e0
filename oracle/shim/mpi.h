/* oracle/shim/mpi.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Single-process stand-in for <mpi.h>, just large enough to compile the reference's
 * Auxiliar_functions.{h,cpp} unmodified (they include <mpi.h> and call MPI_Send/MPI_Recv in
 * removeTileFromRank/addTileToRank, Auxiliar_functions.cpp:242-272).  Point-to-point messages are
 * delivered through an in-process mailbox keyed by tag, so the oracle driver (ref_wrap.cpp) can emulate
 * several ranks inside one process.  It also pulls in <string>, which the reference header needs
 * but forgets (Auxiliar_functions.h:75).
 */
#ifndef ORACLE_SHIM_MPI_H
#define ORACLE_SHIM_MPI_H
#include <string>
#include <cstddef>

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

#ifdef __cplusplus
extern "C" {
#endif
/* implemented in oracle/ref_wrap.cpp */
int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status* st);
#ifdef __cplusplus
}
#endif
#endif
