// comm.cu -- multi-GPU exchange (one process per GPU).  Placeholder for world == 1 until the
// candidate exchange lands; every entry point fails loudly rather than pretending.
#include "dpsvm_internal.cuh"

namespace dpsvm {
void comm_free(dpsvm_ctx*) {}
int comm_launch_iteration(dpsvm_ctx*, cudaStream_t) { return fail(DPSVM_ERR_COMM, "multi-GPU exchange not initialised"); }
}  // namespace dpsvm

extern "C" int dpsvm_comm_unique_id(unsigned char*) { return dpsvm::fail(DPSVM_ERR_COMM, "dpsvm_comm_unique_id: not available in this build"); }
extern "C" int dpsvm_comm_init(dpsvm_ctx*, const unsigned char*, int, int) { return dpsvm::fail(DPSVM_ERR_COMM, "dpsvm_comm_init: not available in this build"); }
