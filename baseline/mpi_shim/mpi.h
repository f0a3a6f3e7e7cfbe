// Stand-in for the handful of MPI entry points the reference uses (SURVEY.md §2.3 lists all
// of them).  The image has no MPI, so the UNMODIFIED reference sources
// (baseline/_ref/engine.cpp, common.cpp) are compiled against this header to obtain the
// reference arm of bench.py.
//
// One rank (the default): collectives copy or do nothing.
// Several ranks: one OS process per rank on one machine, started with RANK / WORLD_SIZE in
// the environment (what torch.distributed.run sets) and DMLP_MPI_DIR naming a directory for
// the rendezvous sockets.  Every pair of ranks is joined by one AF_UNIX stream socket; a
// collective is a fixed sequence of blocking sends and receives between the members of its
// communicator, so the per-pair byte streams stay in step without tags.  Semantics follow
// the MPI standard for exactly the calls and argument patterns of the reference: row-major
// Cartesian ranks, non-increasing balanced MPI_Dims_create, rank-ordered Gather/Scatterv.
// Handles are pointers because the reference passes nullptr as a datatype on non-root ranks
// (SURVEY.md §2.5 B7).
#ifndef DMLP_MPI_SHIM_H_
#define DMLP_MPI_SHIM_H_

#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <sys/types.h>
#include <sys/un.h>
#include <unistd.h>

#include <cerrno>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct shim_type { size_t size; };
struct shim_comm {
  std::vector<int> members;  // world ranks, in communicator rank order
  int me = 0;                // my rank in this communicator
  int dims[2] = {1, 1};      // Cartesian shape (cart communicators)
  bool cart = false;
};
typedef shim_type *MPI_Datatype;
typedef shim_comm *MPI_Comm;
typedef long MPI_Aint;

inline shim_type shim_int_t = {sizeof(int)}, shim_double_t = {sizeof(double)}, shim_byte_t = {1};
inline shim_comm shim_world;
inline std::vector<int> shim_fd;  // socket to every world rank (-1 for myself)
inline int shim_rank = 0, shim_size = 1;
#define MPI_INT (&shim_int_t)
#define MPI_DOUBLE (&shim_double_t)
#define MPI_BYTE (&shim_byte_t)
#define MPI_COMM_WORLD (&shim_world)
#define MPI_SUCCESS 0

inline void shim_die(const char *what) {
  std::fprintf(stderr, "[mpi shim rank %d] %s: %s\n", shim_rank, what, std::strerror(errno));
  std::_Exit(70);
}
inline void shim_send(int peer, const void *buf, size_t bytes) {
  const char *p = static_cast<const char *>(buf);
  while (bytes > 0) {
    ssize_t n = ::send(shim_fd[size_t(peer)], p, bytes, MSG_NOSIGNAL);
    if (n < 0) { if (errno == EINTR) continue; shim_die("send"); }
    p += n; bytes -= size_t(n);
  }
}
inline void shim_recv(int peer, void *buf, size_t bytes) {
  char *p = static_cast<char *>(buf);
  while (bytes > 0) {
    ssize_t n = ::recv(shim_fd[size_t(peer)], p, bytes, 0);
    if (n < 0) { if (errno == EINTR) continue; shim_die("recv"); }
    if (n == 0) { errno = ECONNRESET; shim_die("recv (peer closed)"); }
    p += n; bytes -= size_t(n);
  }
}
inline std::string shim_path(const std::string &dir, int r) { return dir + "/rank_" + std::to_string(r); }

inline int MPI_Init(int *, char ***) {
  const char *rs = std::getenv("RANK"), *ws = std::getenv("WORLD_SIZE"), *dir = std::getenv("DMLP_MPI_DIR");
  shim_rank = rs ? std::atoi(rs) : 0;
  shim_size = (ws && dir) ? std::atoi(ws) : 1;
  if (shim_size < 1) shim_size = 1;
  if (shim_size == 1) shim_rank = 0;
  shim_world.members.resize(size_t(shim_size));
  for (int i = 0; i < shim_size; ++i) shim_world.members[size_t(i)] = i;
  shim_world.me = shim_rank;
  shim_fd.assign(size_t(shim_size), -1);
  if (shim_size == 1) return MPI_SUCCESS;
  // listen first, then connect to the lower ranks and accept the higher ones
  const int ls = ::socket(AF_UNIX, SOCK_STREAM, 0);
  if (ls < 0) shim_die("socket");
  sockaddr_un addr{};
  addr.sun_family = AF_UNIX;
  const std::string mine = shim_path(dir, shim_rank);
  if (mine.size() >= sizeof(addr.sun_path)) { errno = ENAMETOOLONG; shim_die("DMLP_MPI_DIR"); }
  std::strcpy(addr.sun_path, mine.c_str());
  ::unlink(mine.c_str());
  if (::bind(ls, reinterpret_cast<sockaddr *>(&addr), sizeof addr) < 0) shim_die("bind");
  if (::listen(ls, shim_size) < 0) shim_die("listen");
  timeval patience{};  // a rank that never shows up must not hang the others for ever
  patience.tv_sec = 180;
  ::setsockopt(ls, SOL_SOCKET, SO_RCVTIMEO, &patience, sizeof patience);
  for (int peer = 0; peer < shim_rank; ++peer) {
    sockaddr_un pa{};
    pa.sun_family = AF_UNIX;
    std::strcpy(pa.sun_path, shim_path(dir, peer).c_str());
    int fd = -1;
    for (int attempt = 0;; ++attempt) {
      fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
      if (fd < 0) shim_die("socket");
      if (::connect(fd, reinterpret_cast<sockaddr *>(&pa), sizeof pa) == 0) break;
      ::close(fd);
      if (attempt > 30000) shim_die("connect (peer never appeared)");
      ::usleep(2000);
    }
    shim_fd[size_t(peer)] = fd;
    shim_send(peer, &shim_rank, sizeof shim_rank);
  }
  for (int i = shim_rank + 1; i < shim_size; ++i) {
    int fd = ::accept(ls, nullptr, nullptr);
    if (fd < 0) shim_die("accept");
    int who = -1;
    char *p = reinterpret_cast<char *>(&who);
    for (size_t got = 0; got < sizeof who;) {
      ssize_t n = ::recv(fd, p + got, sizeof who - got, 0);
      if (n <= 0) shim_die("recv (hello)");
      got += size_t(n);
    }
    if (who <= shim_rank || who >= shim_size || shim_fd[size_t(who)] != -1) { errno = EPROTO; shim_die("hello"); }
    shim_fd[size_t(who)] = fd;
  }
  ::close(ls);
  ::unlink(mine.c_str());
  return MPI_SUCCESS;
}
inline int MPI_Comm_rank(MPI_Comm c, int *rank) { *rank = c->me; return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm c, int *size) { *size = c->members.empty() ? 1 : int(c->members.size()); return MPI_SUCCESS; }

inline int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm c) {
  const int n = int(c->members.size());
  if (n <= 1) return MPI_SUCCESS;
  const size_t bytes = size_t(count) * type->size;
  if (c->me == root) {
    for (int r = 0; r < n; ++r)
      if (r != root) shim_send(c->members[size_t(r)], buf, bytes);
  } else {
    shim_recv(c->members[size_t(root)], buf, bytes);
  }
  return MPI_SUCCESS;
}
inline int MPI_Gather(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, int recvcount,
                      MPI_Datatype recvtype, int root, MPI_Comm c) {
  const int n = c->members.empty() ? 1 : int(c->members.size());
  const size_t sbytes = size_t(sendcount) * (sendtype ? sendtype->size : 0);
  if (c->me != root) {
    shim_send(c->members[size_t(root)], sendbuf, sbytes);
    return MPI_SUCCESS;
  }
  const size_t rbytes = size_t(recvcount) * (recvtype ? recvtype->size : 0);
  for (int r = 0; r < n; ++r) {
    char *dst = static_cast<char *>(recvbuf) + size_t(r) * rbytes;
    if (r == root) { if (sbytes) std::memcpy(dst, sendbuf, sbytes); }
    else shim_recv(c->members[size_t(r)], dst, rbytes);
  }
  return MPI_SUCCESS;
}
inline int MPI_Scatterv(const void *sendbuf, const int *sendcounts, const int *displs, MPI_Datatype sendtype,
                        void *recvbuf, int recvcount, MPI_Datatype recvtype, int root, MPI_Comm c) {
  const int n = c->members.empty() ? 1 : int(c->members.size());
  if (c->me != root) {
    shim_recv(c->members[size_t(root)], recvbuf, size_t(recvcount) * recvtype->size);
    return MPI_SUCCESS;
  }
  for (int r = 0; r < n; ++r) {
    const char *src = static_cast<const char *>(sendbuf) + size_t(displs[r]) * sendtype->size;
    const size_t bytes = size_t(sendcounts[r]) * sendtype->size;
    if (r == root) { if (bytes && recvbuf) std::memcpy(recvbuf, src, bytes); }
    else shim_send(c->members[size_t(r)], src, bytes);
  }
  return MPI_SUCCESS;
}
inline int MPI_Barrier(MPI_Comm c) {
  const int n = int(c->members.size());
  if (n <= 1) return MPI_SUCCESS;
  char token = 0;
  if (c->me == 0) {
    for (int r = 1; r < n; ++r) shim_recv(c->members[size_t(r)], &token, 1);
    for (int r = 1; r < n; ++r) shim_send(c->members[size_t(r)], &token, 1);
  } else {
    shim_send(c->members[0], &token, 1);
    shim_recv(c->members[0], &token, 1);
  }
  return MPI_SUCCESS;
}
inline int MPI_Finalize() {
  if (shim_size > 1) MPI_Barrier(&shim_world);
  for (int fd : shim_fd) if (fd >= 0) ::close(fd);
  shim_fd.clear();
  return MPI_SUCCESS;
}

// Balanced factorisation in non-increasing order, as MPI_Dims_create specifies for free (zero)
// entries: 2 -> 2x1, 4 -> 2x2, 6 -> 3x2, 8 -> 4x2, 12 -> 4x3.
inline int MPI_Dims_create(int nnodes, int ndims, int *dims) {
  for (int i = 0; i < ndims; ++i) dims[i] = 1;
  if (ndims < 2) { if (ndims == 1) dims[0] = nnodes; return MPI_SUCCESS; }
  int best = 1;
  for (int f = 1; f * f <= nnodes; ++f)
    if (nnodes % f == 0) best = f;
  dims[0] = nnodes / best;
  dims[1] = best;
  return MPI_SUCCESS;
}
inline int MPI_Cart_create(MPI_Comm c, int ndims, const int *dims, const int *, int, MPI_Comm *out) {
  shim_comm *n = new shim_comm(*c);
  n->cart = true;
  n->dims[0] = ndims > 0 ? dims[0] : 1;
  n->dims[1] = ndims > 1 ? dims[1] : 1;
  *out = n;
  return MPI_SUCCESS;
}
inline int MPI_Cart_coords(MPI_Comm c, int rank, int maxdims, int *coords) {
  if (maxdims > 0) coords[0] = rank / c->dims[1];
  if (maxdims > 1) coords[1] = rank % c->dims[1];
  for (int i = 2; i < maxdims; ++i) coords[i] = 0;
  return MPI_SUCCESS;
}
// Keep the dimensions flagged in remain_dims: the new communicator holds the ranks that
// share my coordinates in the dropped dimensions, ordered by the kept coordinate.
inline int MPI_Cart_sub(MPI_Comm c, const int *remain_dims, MPI_Comm *out) {
  shim_comm *n = new shim_comm;
  const int my_r = c->me / c->dims[1], my_c = c->me % c->dims[1];
  for (int r = 0; r < int(c->members.size()); ++r) {
    const int rr = r / c->dims[1], rc = r % c->dims[1];
    const bool same = (remain_dims[0] || rr == my_r) && (remain_dims[1] || rc == my_c);
    if (!same) continue;
    if (r == c->me) n->me = int(n->members.size());
    n->members.push_back(c->members[size_t(r)]);
  }
  n->cart = true;
  n->dims[0] = remain_dims[0] ? c->dims[0] : 1;
  n->dims[1] = remain_dims[1] ? c->dims[1] : 1;
  *out = n;
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
