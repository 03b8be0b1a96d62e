// TEST INFRASTRUCTURE — NCCL over files for the CUDA-on-CPU emulation (see nccl.h in this directory)
#include "nccl.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>
#include <unistd.h>

struct emu_nccl_comm {
    std::string dir; int n = 1, rank = 0;
    std::map<int, long> sseq, rseq; long arseq = 0;
};
namespace {
struct Pending { void* buf; size_t bytes; int peer; ncclComm_t comm; };
std::vector<Pending> pending;
int group_depth = 0;
bool read_file(const std::string& path, void* buf, size_t bytes, bool unlink_after) {
    for (int tries = 0; tries < 120000; tries++) {                      // up to ~2 minutes
        FILE* f = fopen(path.c_str(), "rb");
        if (f) {
            const size_t got = fread(buf, 1, bytes, f);
            fclose(f);
            if (got != bytes) return false;
            if (unlink_after) unlink(path.c_str());
            return true;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    return false;
}
bool write_file(const std::string& path, const void* buf, size_t bytes) {
    const std::string tmp = path + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return false;
    const size_t put = fwrite(buf, 1, bytes, f);
    fclose(f);
    return put == bytes && rename(tmp.c_str(), path.c_str()) == 0;       // rename: the reader never sees a partial file
}
ncclResult_t do_recv(const Pending& p) {
    emu_nccl_comm& c = *p.comm;
    const std::string path = c.dir + "/p2p_s" + std::to_string(p.peer) + "_d" + std::to_string(c.rank) + "_" + std::to_string(c.rseq[p.peer]++) + ".bin";
    return read_file(path, p.buf, p.bytes, true) ? 0 : 1;
}
}  // namespace

const char* ncclGetErrorString(ncclResult_t r) { return r ? "emulated NCCL: peer data did not arrive" : "ok"; }
ncclResult_t ncclGetUniqueId(ncclUniqueId* id) {
    memset(id->internal, 0, sizeof(id->internal));
    char tmpl[] = "/tmp/fb200_emu_nccl_XXXXXX";
    if (!mkdtemp(tmpl)) return 1;
    strncpy(id->internal, tmpl, sizeof(id->internal) - 1);
    return 0;
}
ncclResult_t ncclCommInitRank(ncclComm_t* comm, int nranks, ncclUniqueId id, int rank) {
    emu_nccl_comm* c = new emu_nccl_comm();
    c->dir = id.internal; c->n = nranks; c->rank = rank;
    *comm = c;
    return 0;
}
ncclResult_t ncclCommDestroy(ncclComm_t comm) {
    if (comm->rank == 0) {                      // best effort: all-reduce files stay behind, p2p files are removed by their readers
        for (long s = 0; s < comm->arseq; s++) for (int r = 0; r < comm->n; r++) unlink((comm->dir + "/ar" + std::to_string(s) + "_r" + std::to_string(r) + ".bin").c_str());
        rmdir(comm->dir.c_str());
    }
    delete comm;
    return 0;
}
ncclResult_t ncclGroupStart() { group_depth++; return 0; }
ncclResult_t ncclGroupEnd() {
    if (--group_depth > 0) return 0;
    ncclResult_t r = 0;
    for (const Pending& p : pending) r |= do_recv(p);
    pending.clear();
    return r;
}
ncclResult_t ncclSend(const void* buf, size_t count, ncclDataType_t dt, int peer, ncclComm_t comm, void*) {
    emu_nccl_comm& c = *comm;
    const std::string path = c.dir + "/p2p_s" + std::to_string(c.rank) + "_d" + std::to_string(peer) + "_" + std::to_string(c.sseq[peer]++) + ".bin";
    return write_file(path, buf, count * (size_t)dt) ? 0 : 1;
}
ncclResult_t ncclRecv(void* buf, size_t count, ncclDataType_t dt, int peer, ncclComm_t comm, void*) {
    Pending p{buf, count * (size_t)dt, peer, comm};
    if (group_depth > 0) { pending.push_back(p); return 0; }
    return do_recv(p);
}
ncclResult_t ncclAllGather(const void* send, void* recv, size_t sendcount, ncclDataType_t dt, ncclComm_t comm, void*) {
    emu_nccl_comm& c = *comm;
    const long s = c.arseq++;
    const size_t bytes = sendcount * (size_t)dt;
    if (!write_file(c.dir + "/ar" + std::to_string(s) + "_r" + std::to_string(c.rank) + ".bin", send, bytes)) return 1;
    for (int r = 0; r < c.n; r++)
        if (!read_file(c.dir + "/ar" + std::to_string(s) + "_r" + std::to_string(r) + ".bin", (char*)recv + (size_t)r * bytes, bytes, false)) return 1;
    return 0;
}
ncclResult_t ncclAllReduce(const void* send, void* recv, size_t count, ncclDataType_t, ncclRedOp_t, ncclComm_t comm, void*) {
    emu_nccl_comm& c = *comm;
    const long s = c.arseq++;
    if (!write_file(c.dir + "/ar" + std::to_string(s) + "_r" + std::to_string(c.rank) + ".bin", send, count * sizeof(double))) return 1;
    std::vector<double> acc(count, 0.0), part(count);
    for (int r = 0; r < c.n; r++) {
        if (!read_file(c.dir + "/ar" + std::to_string(s) + "_r" + std::to_string(r) + ".bin", part.data(), count * sizeof(double), false)) return 1;
        for (size_t i = 0; i < count; i++) acc[i] = r == 0 ? part[i] : acc[i] + part[i];
    }
    memcpy(recv, acc.data(), count * sizeof(double));
    return 0;
}
ncclResult_t ncclBroadcast(const void* send, void* recv, size_t count, ncclDataType_t dt, int root, ncclComm_t comm, void*) {
    emu_nccl_comm& c = *comm;
    const long s = c.arseq++;
    const size_t bytes = count * (size_t)dt;
    const std::string path = c.dir + "/ar" + std::to_string(s) + "_r" + std::to_string(root) + ".bin";
    if (c.rank == root) {
        if (!write_file(path, send, bytes)) return 1;
        if (recv != send) memcpy(recv, send, bytes);
        return 0;
    }
    return read_file(path, recv, bytes, false) ? 0 : 1;
}
