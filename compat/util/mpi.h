// Minimal file-based stand-in for the handful of MPI calls the reference ring driver makes
// (MPI_Init/Comm_rank/Comm_size/Bcast/Barrier/Gather/Finalize).  Ranks are separate
// processes started with RANK / WORLD_SIZE in the environment; data moves through files in
// $B200_MPI_DIR (default /tmp/b200_mpi).  Test tooling for the drop-in check, not product code.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unistd.h>
#include <sys/stat.h>
typedef int MPI_Comm; typedef int MPI_Datatype;
#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_FLOAT 4
#define MPI_BYTE 1
namespace b200mpi {
inline int& rank() { static int r = getenv("RANK") ? atoi(getenv("RANK")) : 0; return r; }
inline int& size() { static int s = getenv("WORLD_SIZE") ? atoi(getenv("WORLD_SIZE")) : 1; return s; }
inline std::string dir() { const char* d = getenv("B200_MPI_DIR"); std::string s = d ? d : "/tmp/b200_mpi"; mkdir(s.c_str(), 0777); return s; }
inline int& seq() { static int s = 0; return s; }
inline void put(const std::string& name, const void* p, size_t n) {
    std::string tmp = dir() + "/" + name + ".tmp", fin = dir() + "/" + name;
    FILE* f = fopen(tmp.c_str(), "wb"); fwrite(p, 1, n, f); fclose(f); rename(tmp.c_str(), fin.c_str());
}
inline void get(const std::string& name, void* p, size_t n) {
    std::string fin = dir() + "/" + name;
    for (;;) { FILE* f = fopen(fin.c_str(), "rb"); if (f) { size_t g = fread(p, 1, n, f); fclose(f); if (g == n) return; } usleep(2000); }
}
}
inline int MPI_Init(int*, char***) { return 0; }
inline int MPI_Finalize() { return 0; }
inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = b200mpi::rank(); return 0; }
inline int MPI_Comm_size(MPI_Comm, int* s) { *s = b200mpi::size(); return 0; }
inline int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm) {
    if (b200mpi::size() == 1) return 0;
    std::string n = "bcast" + std::to_string(b200mpi::seq()++);
    if (b200mpi::rank() == root) b200mpi::put(n, buf, (size_t)count * t); else b200mpi::get(n, buf, (size_t)count * t);
    return 0;
}
inline int MPI_Barrier(MPI_Comm) {
    if (b200mpi::size() == 1) return 0;
    std::string n = "bar" + std::to_string(b200mpi::seq()++);
    char c = 1; b200mpi::put(n + "_" + std::to_string(b200mpi::rank()), &c, 1);
    for (int r = 0; r < b200mpi::size(); ++r) b200mpi::get(n + "_" + std::to_string(r), &c, 1);
    return 0;
}
inline int MPI_Gather(const void* s, int sc, MPI_Datatype st, void* r, int rc, MPI_Datatype rt, int root, MPI_Comm) {
    std::string n = "gather" + std::to_string(b200mpi::seq()++);
    if (b200mpi::size() == 1) { memcpy(r, s, (size_t)sc * st); return 0; }
    b200mpi::put(n + "_" + std::to_string(b200mpi::rank()), s, (size_t)sc * st);
    if (b200mpi::rank() == root)
        for (int k = 0; k < b200mpi::size(); ++k) b200mpi::get(n + "_" + std::to_string(k), (char*)r + (size_t)k * rc * rt, (size_t)rc * rt);
    return 0;
}
// 128-byte NCCL id broadcast used by init_nccl_comm
extern "C" int b200_ring_unique_id(void*);
inline int MPI_Bcast_id(unsigned char* id, int rank) {
    if (rank == 0 && b200_ring_unique_id(id) != 0) return -1;
    return MPI_Bcast(id, 128, MPI_BYTE, 0, MPI_COMM_WORLD);
}
