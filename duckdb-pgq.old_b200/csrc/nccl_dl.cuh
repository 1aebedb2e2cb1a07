// NCCL entry points resolved with dlopen at first use, so libpgq_b200.so loads (and the
// single-GPU path runs) on hosts without NCCL.  Only the 1-D vertex-partitioned path
// (BASELINE config 5, part.cu) needs a communicator.
#pragma once
#include <nccl.h>

namespace pgq {

struct NcclApi {
	ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
	                          cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char *(*GetErrorString)(ncclResult_t) = nullptr;
	bool ok = false;
};

// nullptr (and pgq_last_error set) when libnccl.so.2 cannot be loaded
const NcclApi *nccl_api();

#define PGQ_NCCL(api, call)                                                                          \
	do {                                                                                             \
		ncclResult_t _r = (call);                                                                    \
		if (_r != ncclSuccess) {                                                                     \
			pgq::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, (api)->GetErrorString(_r)); \
			return PGQ_ERR_CUDA;                                                                     \
		}                                                                                            \
	} while (0)

} // namespace pgq
