// Library-level pieces of the C ABI: version, error string, argument validation.
#include <stdlib.h>
#include <string.h>

#include "k7_common.cuh"

static thread_local char g_err[512] = "";

void k7_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* k7_last_error(void) { return g_err; }
extern "C" int k7_version(void) { return K7_ABI_VERSION; }

static unsigned long long g_launches = 0;
void k7_count_launch() { __atomic_fetch_add(&g_launches, 1ull, __ATOMIC_RELAXED); }
extern "C" unsigned long long k7_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int k7_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

int k7_check_spec(const K7AnchorSpec* s, int A) {
    K7_REQUIRE(s != nullptr, "anchor spec is null");
    K7_REQUIRE(s->num_levels >= 1 && s->num_levels <= K7_MAX_LEVELS, "anchor spec: num_levels=%d out of range", s->num_levels);
    K7_REQUIRE(s->cell_anchors >= 1 && s->cell_anchors <= K7_MAX_CELL_ANCHORS, "anchor spec: cell_anchors=%d out of range",
               s->cell_anchors);
    K7_REQUIRE(s->box_order == K7_BOX_XYXY || s->box_order == K7_BOX_YXYX, "anchor spec: bad box_order");
    long long off = 0;
    for (int l = 0; l < s->num_levels; ++l) {
        K7_REQUIRE(s->fh[l] >= 0 && s->fw[l] >= 0, "anchor spec: negative grid at level %d", l);
        K7_REQUIRE(s->offset[l] == off, "anchor spec: offset[%d]=%d, expected %lld", l, s->offset[l], off);
        off += (long long)s->fh[l] * s->fw[l] * s->cell_anchors;
    }
    K7_REQUIRE(off == s->total && s->offset[s->num_levels] == s->total, "anchor spec: total=%d, levels sum to %lld",
               s->total, off);
    K7_REQUIRE(s->total == A, "anchor spec describes %d anchors but the tensors have A=%d", s->total, A);
    return K7_OK;
}


// ------------------------------------------------------------------ peer mailboxes (K7PeerReduce)
// The NVLink plumbing of the fused loss-sum exchange: a mailbox is its own cudaMalloc allocation on
// the caller's current device so that it can be exported with the CUDA IPC API; the host exchanges
// the 64-byte handles with whatever transport it has (torch.distributed in b200det) and opens the
// peers' handles, which also enables peer access from the current device.
static_assert(sizeof(cudaIpcMemHandle_t) == K7_IPC_HANDLE_BYTES, "K7_IPC_HANDLE_BYTES");

extern "C" int k7_peer_mailbox_create(int world, float** mailbox, unsigned char* handle) {
    K7_REQUIRE(mailbox && handle, "k7_peer_mailbox_create: null argument");
    K7_REQUIRE(world >= 1 && world <= K7_MAX_PEERS, "k7_peer_mailbox_create: world=%d out of range", world);
    void* p = nullptr;
    const size_t bytes = sizeof(float) * K7_PEER_MAILBOX_FLOATS(world);
    K7_CUDA(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        k7_set_error("k7_peer_mailbox_create: %s", cudaGetErrorString(e));
        return K7_ECUDA;
    }
    memcpy(handle, &h, sizeof(h));
    *mailbox = static_cast<float*>(p);
    return K7_OK;
}

extern "C" int k7_peer_mailbox_open(const unsigned char* handle, float** mailbox) {
    K7_REQUIRE(mailbox && handle, "k7_peer_mailbox_open: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    K7_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *mailbox = static_cast<float*>(p);
    return K7_OK;
}

extern "C" int k7_peer_mailbox_close(float* mailbox) {
    if (mailbox) K7_CUDA(cudaIpcCloseMemHandle(mailbox));
    return K7_OK;
}

extern "C" int k7_peer_mailbox_destroy(float* mailbox) {
    if (mailbox) K7_CUDA(cudaFree(mailbox));
    return K7_OK;
}
