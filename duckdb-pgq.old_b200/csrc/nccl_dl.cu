#include <dlfcn.h>

#include <mutex>

#include "common.cuh"
#include "nccl_dl.cuh"

namespace pgq {

const NcclApi *nccl_api() {
	static NcclApi api;
	static std::once_flag once;
	std::call_once(once, [] {
		// a process that already loaded a libnccl.so.2 (e.g. torch's bundled one) gets that copy back
		void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
		if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
		if (!h) return;
		bool all = true;
		auto sym = [&](const char *name) {
			void *p = dlsym(h, name);
			if (!p) all = false;
			return p;
		};
		api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
		api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
		api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
		api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
		api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
		api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
		api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
		api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
		api.ok = all;
	});
	if (!api.ok) {
		set_error("NCCL (libnccl.so.2) is not available: %s", dlerror() ? dlerror() : "missing symbols");
		return nullptr;
	}
	return &api;
}

} // namespace pgq
