#include "nccl_dyn.h"

#include <dlfcn.h>

#include <mutex>

namespace mnenc {

const NcclApi* ncclApi(std::string* err) {
    static NcclApi api{};
    static std::string failure;
    static bool ok = false;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) {
            failure = std::string("cannot load libnccl.so.2: ") + dlerror();
            return;
        }
        bool all = true;
        auto sym = [&](const char* name) {
            void* p = dlsym(h, name);
            if (!p) {
                all = false;
                failure = std::string("libnccl lacks ") + name;
            }
            return p;
        };
        api.getUniqueId = reinterpret_cast<decltype(api.getUniqueId)>(sym("ncclGetUniqueId"));
        api.commInitRank = reinterpret_cast<decltype(api.commInitRank)>(sym("ncclCommInitRank"));
        api.commInitAll = reinterpret_cast<decltype(api.commInitAll)>(sym("ncclCommInitAll"));
        api.commDestroy = reinterpret_cast<decltype(api.commDestroy)>(sym("ncclCommDestroy"));
        api.allGather = reinterpret_cast<decltype(api.allGather)>(sym("ncclAllGather"));
        api.allReduce = reinterpret_cast<decltype(api.allReduce)>(sym("ncclAllReduce"));
        api.reduce = reinterpret_cast<decltype(api.reduce)>(sym("ncclReduce"));
        api.send = reinterpret_cast<decltype(api.send)>(sym("ncclSend"));
        api.recv = reinterpret_cast<decltype(api.recv)>(sym("ncclRecv"));
        api.groupStart = reinterpret_cast<decltype(api.groupStart)>(sym("ncclGroupStart"));
        api.groupEnd = reinterpret_cast<decltype(api.groupEnd)>(sym("ncclGroupEnd"));
        api.getErrorString = reinterpret_cast<decltype(api.getErrorString)>(sym("ncclGetErrorString"));
        ok = all;
    });
    if (!ok) {
        if (err) *err = failure;
        return nullptr;
    }
    return &api;
}

}  // namespace mnenc
