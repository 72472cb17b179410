/* Minimal stand-in for <mpi.h>, for building reference callers (examples/wordfreq.cpp, examples/rmat.cpp,
 * cpu/IntCount.cpp ...) against this library on a box without MPI -- the role the reference's own mpistubs/
 * plays (mpistubs/mpi.h:19-125).  Written from scratch; only what those callers use.
 * Unlike mpistubs it is not limited to one process: rank and size come from the environment a process launcher
 * sets (RANK / WORLD_SIZE as torch.distributed.run does, or OMPI_COMM_WORLD_* / PMI_*), one process per GPU, and
 * MapReduce(MPI_Comm) hands the 128-byte NCCL id from rank 0 to the others over a loopback TCP socket
 * (cpp/mapreduce.cpp, MRGPU_PORT).  Data never goes through this header: every exchange is NCCL inside the library.
 * With a real MPI, use its mpi.h instead: MapReduce(MPI_Comm) then bootstraps NCCL with one MPI_Bcast of 128 bytes
 * (INTEGRATION.md section 1). */
#ifndef MRGPU_MPI_STUB_H
#define MRGPU_MPI_STUB_H

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

typedef int MPI_Comm;
#define MPI_COMM_WORLD 0
#define MPI_MAX_PROCESSOR_NAME 128
#define MRGPU_MPI_STUB 1

static inline int MPI_Init(int *argc, char ***argv) { return 0; }
static inline int MPI_Initialized(int *flag)
{
  *flag = 1;
  return 0;
}
static inline int MPI_Finalize(void) { return 0; }
static inline int mrgpu_stub_env(const char *a, const char *b, const char *c, int dflt)
{
  const char *names[3] = {a, b, c};
  for (int i = 0; i < 3; i++) {
    const char *v = names[i] ? getenv(names[i]) : NULL;
    if (v && *v) return atoi(v);
  }
  return dflt;
}
static inline int MPI_Comm_rank(MPI_Comm comm, int *me)
{
  *me = mrgpu_stub_env("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", 0);
  return 0;
}
static inline int MPI_Comm_size(MPI_Comm comm, int *n)
{
  *n = mrgpu_stub_env("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", 1);
  return 0;
}
static inline int MPI_Barrier(MPI_Comm comm) { return 0; }

// Host-side collectives, as far as OINK's front-end uses them (oink/*.cpp: MPI_Bcast of input lines, MPI_Allreduce of
// counters, MPI_Comm_split for "universe" partitions).  With one process they are copies, as in mpistubs/mpi.cpp.
// With several processes this header cannot move host data between them -- only the library moves data (NCCL) -- so
// they stop the program instead of returning a wrong answer: a multi-rank OINK needs a real MPI.
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_INT 1
#define MPI_CHAR 2
#define MPI_BYTE 3
#define MPI_UNSIGNED_LONG 4
#define MPI_DOUBLE 5
#define MPI_FLOAT 6
#define MPI_LONG 7
#define MPI_UNSIGNED_LONG_LONG 8
#define MPI_SUM 1
#define MPI_MIN 2
#define MPI_MAX 3
static inline size_t mrgpu_stub_type_size(MPI_Datatype t)
{
  switch (t) {
    case MPI_INT: case MPI_FLOAT: return 4;
    case MPI_CHAR: case MPI_BYTE: return 1;
    default: return 8;
  }
}
static inline void mrgpu_stub_single_process_only(const char *what)
{
  if (mrgpu_stub_env("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", 1) > 1) {
    printf("ERROR: %s among %s needs a real MPI (cpp/mpi.h is a stand-in that moves no host data)\n", what, "several processes");
    fflush(stdout);
    exit(1);
  }
}
static inline int MPI_Allreduce(void *in, void *out, int n, MPI_Datatype t, MPI_Op op, MPI_Comm comm)
{
  mrgpu_stub_single_process_only("MPI_Allreduce");
  if (in != out) memcpy(out, in, (size_t) n * mrgpu_stub_type_size(t));
  return 0;
}
static inline int MPI_Bcast(void *buf, int n, MPI_Datatype t, int root, MPI_Comm comm)
{
  mrgpu_stub_single_process_only("MPI_Bcast");
  return 0;
}
static inline int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *out)
{
  mrgpu_stub_single_process_only("MPI_Comm_split");
  *out = comm;
  return 0;
}
static inline int MPI_Comm_free(MPI_Comm *comm) { return 0; }
static inline int MPI_Abort(MPI_Comm comm, int code)
{
  exit(code);
  return 0;
}
static inline double MPI_Wtime(void)
{
  struct timeval tv;
  gettimeofday(&tv, NULL);
  return tv.tv_sec + 1.0e-6 * tv.tv_usec;
}
static inline int MPI_Get_processor_name(char *name, int *len)
{
  strcpy(name, "gpu");
  *len = 3;
  return 0;
}

#endif
