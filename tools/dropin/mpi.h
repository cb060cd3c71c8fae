// Single-rank stand-in for the OpenMPI C++ bindings the reference driver uses
// (reference svmTrainMain.cpp:144,153,154,159,183,191,198,233,244,362), so that the UNMODIFIED
// reference main() can be linked against this repo's SvmTrain as a drop-in proof.  Test tooling only.
#ifndef DPSVM_DROPIN_MPI_STUB_H
#define DPSVM_DROPIN_MPI_STUB_H
#include <string.h>
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_FLOAT 4
namespace MPI {
inline void Init(int&, char**&) {}
inline void Finalize() {}
inline void Get_processor_name(char* name, int& len) { strcpy(name, "localhost"); len = 9; }
struct Comm {
    int Get_size() const { return 1; }
    int Get_rank() const { return 0; }
    void Barrier() const {}
    void Allgather(const void* send, int count, int elem_bytes, void* recv, int, int) const
    {
        memcpy(recv, send, (size_t)count * (size_t)elem_bytes);
    }
};
static Comm COMM_WORLD;
}  // namespace MPI
#endif
