// Single-process stand-in for the handful of MPI entry points the reference uses
// (SURVEY.md §2.3 lists all of them).  The image has no MPI, so the UNMODIFIED
// reference sources (baseline/_ref/engine.cpp, common.cpp) are compiled against
// this header to obtain the reference arm of bench.py at one rank.  Everything
// is the degenerate world-size-1 case: collectives copy or do nothing.
// Handles are pointers because the reference passes nullptr as a datatype on
// non-root ranks.
#ifndef DMLP_MPI_SHIM_H_
#define DMLP_MPI_SHIM_H_

#include <cstddef>
#include <cstring>

struct shim_type { size_t size; };
struct shim_comm { int size; };
typedef shim_type *MPI_Datatype;
typedef shim_comm *MPI_Comm;
typedef long MPI_Aint;

static shim_type shim_int_t = {sizeof(int)}, shim_double_t = {sizeof(double)}, shim_byte_t = {1};
static shim_comm shim_world = {1};
#define MPI_INT (&shim_int_t)
#define MPI_DOUBLE (&shim_double_t)
#define MPI_BYTE (&shim_byte_t)
#define MPI_COMM_WORLD (&shim_world)
#define MPI_SUCCESS 0

inline int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
inline int MPI_Finalize() { return MPI_SUCCESS; }
inline int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = 0; return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm, int *size) { *size = 1; return MPI_SUCCESS; }
inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
inline int MPI_Bcast(void *, int, MPI_Datatype, int, MPI_Comm) { return MPI_SUCCESS; }
inline int MPI_Dims_create(int nnodes, int ndims, int *dims) {
  for (int i = 0; i < ndims; ++i) dims[i] = 1;
  dims[0] = nnodes;
  return MPI_SUCCESS;
}
inline int MPI_Cart_create(MPI_Comm, int, const int *, const int *, int, MPI_Comm *out) {
  *out = &shim_world;
  return MPI_SUCCESS;
}
inline int MPI_Cart_coords(MPI_Comm, int, int maxdims, int *coords) {
  for (int i = 0; i < maxdims; ++i) coords[i] = 0;
  return MPI_SUCCESS;
}
inline int MPI_Cart_sub(MPI_Comm, const int *, MPI_Comm *out) { *out = &shim_world; return MPI_SUCCESS; }
inline int MPI_Gather(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, int, MPI_Datatype,
                      int, MPI_Comm) {
  if (recvbuf && sendbuf && sendtype) std::memcpy(recvbuf, sendbuf, size_t(sendcount) * sendtype->size);
  return MPI_SUCCESS;
}
inline int MPI_Scatterv(const void *sendbuf, const int *sendcounts, const int *displs, MPI_Datatype sendtype,
                        void *recvbuf, int, MPI_Datatype, int, MPI_Comm) {
  if (sendbuf && recvbuf && sendtype)
    std::memcpy(recvbuf, static_cast<const char *>(sendbuf) + size_t(displs[0]) * sendtype->size,
                size_t(sendcounts[0]) * sendtype->size);
  return MPI_SUCCESS;
}
inline int MPI_Get_address(const void *location, MPI_Aint *address) {
  *address = reinterpret_cast<MPI_Aint>(location);
  return MPI_SUCCESS;
}
inline int MPI_Type_create_struct(int count, const int *blocklengths, const MPI_Aint *displs,
                                  const MPI_Datatype *types, MPI_Datatype *newtype) {
  size_t extent = 0;
  for (int i = 0; i < count; ++i) {
    size_t end = size_t(displs[i]) + size_t(blocklengths[i]) * types[i]->size;
    if (end > extent) extent = end;
  }
  extent = (extent + 7) / 8 * 8;  // the reference's {double, int, int} tuple: 16 bytes
  *newtype = new shim_type{extent};
  return MPI_SUCCESS;
}
inline int MPI_Type_commit(MPI_Datatype *) { return MPI_SUCCESS; }
inline int MPI_Type_free(MPI_Datatype *) { return MPI_SUCCESS; }

#endif  // DMLP_MPI_SHIM_H_
