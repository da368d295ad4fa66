// Part of ggrm_api.cu: device staging slots for hosts that feed SEVERAL handles from ONE host read of the BED.
// 2D-tile multi-GPU mode: every GPU accumulates every SNP.  Instead of each GPU pulling the whole file over its own PCIe link
// (and through its own pinned ring: C5 = 30 GB per GPU), the host uploads chunk j once, to the GPU j % N, and every handle
// takes it from that GPU's staging slot with ggrm_push_device — a peer copy over NVLink (input side of the 2D-tile
// partition).

namespace ggrm {

struct Stage {
    int device = 0;
    int64_t slot_bytes = 0;
    std::vector<uint8_t*> slots;
    cudaStream_t stream = nullptr;
    static constexpr int kBounce = 2;
    static constexpr int64_t kBounceBytes = int64_t{32} << 20;
    uint8_t* bounce[kBounce] = {nullptr, nullptr};   // pinned; pageable sources (the mapped .bed) are staged through them
    cudaEvent_t bounce_free[kBounce] = {nullptr, nullptr};
    std::string last_error;
};

static void stage_free(Stage* s) {
    if (s == nullptr) return;
    cudaSetDevice(s->device);
    if (s->stream != nullptr) cudaStreamSynchronize(s->stream);
    for (uint8_t* p : s->slots) if (p != nullptr) cudaFree(p);
    for (int i = 0; i < Stage::kBounce; ++i) {
        if (s->bounce[i] != nullptr) cudaFreeHost(s->bounce[i]);
        if (s->bounce_free[i] != nullptr) cudaEventDestroy(s->bounce_free[i]);
    }
    if (s->stream != nullptr) cudaStreamDestroy(s->stream);
    delete s;
}

}  // namespace ggrm

#define GGRM_STAGE_BEGIN(s)                                 \
    if ((s) == nullptr) return GGRM_ERR_INVALID_ARGUMENT;   \
    try {                                                   \
        GGRM_CUDA(cudaSetDevice((s)->device));
#define GGRM_STAGE_END(s)                                   \
    }                                                       \
    catch (const CudaError& e) {                            \
        (s)->last_error = e.msg;                            \
        g_create_error = e.msg;                             \
        return e.status;                                    \
    }                                                       \
    catch (const std::exception& e) {                       \
        (s)->last_error = e.what();                         \
        g_create_error = e.what();                          \
        return GGRM_ERR_INVALID_OPERATION;                  \
    }                                                       \
    return GGRM_OK;

extern "C" int ggrm_stage_create(int device, int64_t slot_bytes, int n_slots, ggrm_stage_t** out) {
    if (out == nullptr) {
        g_create_error = "null argument";
        return GGRM_ERR_INVALID_ARGUMENT;
    }
    *out = nullptr;
    Stage* s = nullptr;
    try {
        if (slot_bytes <= 0 || n_slots < 1 || n_slots > 64) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "stage: slot_bytes > 0 and 1 <= n_slots <= 64 required"};
        check_device(device, nullptr);
        s = new Stage();
        s->device = device;
        s->slot_bytes = slot_bytes;
        GGRM_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
        for (int i = 0; i < n_slots; ++i) {
            void* p = nullptr;
            GGRM_CUDA(cudaMalloc(&p, static_cast<size_t>(slot_bytes)));
            s->slots.push_back(static_cast<uint8_t*>(p));
        }
    } catch (const CudaError& e) {
        g_create_error = e.msg;
        stage_free(s);
        return e.status;
    } catch (const std::exception& e) {
        g_create_error = e.what();
        stage_free(s);
        return GGRM_ERR_OUT_OF_MEMORY;
    }
    *out = reinterpret_cast<ggrm_stage_t*>(s);
    return GGRM_OK;
}

extern "C" void ggrm_stage_destroy(ggrm_stage_t* ss) { stage_free(reinterpret_cast<Stage*>(ss)); }

extern "C" int ggrm_stage_upload(ggrm_stage_t* ss, int slot, const uint8_t* host_src, int64_t bytes) {
    Stage* s = reinterpret_cast<Stage*>(ss);
    GGRM_STAGE_BEGIN(s)
    if (slot < 0 || slot >= static_cast<int>(s->slots.size())) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "stage: slot out of range"};
    if (bytes < 0 || bytes > s->slot_bytes) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "stage: upload larger than the slot"};
    if (host_src == nullptr && bytes > 0) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "null payload"};
    uint8_t* dst = s->slots[static_cast<size_t>(slot)];
    bool pinned = false;
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, host_src) == cudaSuccess) pinned = attr.type == cudaMemoryTypeHost;
    else cudaGetLastError();
    if (pinned) {
        GGRM_CUDA(cudaMemcpyAsync(dst, host_src, static_cast<size_t>(bytes), cudaMemcpyHostToDevice, s->stream));
    } else {
        if (s->bounce[0] == nullptr) {
            for (int i = 0; i < Stage::kBounce; ++i) {
                GGRM_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&s->bounce[i]), static_cast<size_t>(Stage::kBounceBytes), cudaHostAllocDefault));
                GGRM_CUDA(cudaEventCreateWithFlags(&s->bounce_free[i], cudaEventDisableTiming));
            }
        }
        int which = 0;
        for (int64_t done = 0; done < bytes; which ^= 1) {
            const int64_t n = std::min<int64_t>(Stage::kBounceBytes, bytes - done);
            GGRM_CUDA(cudaEventSynchronize(s->bounce_free[which]));  // the DMA that last read this bounce buffer has finished
            std::memcpy(s->bounce[which], host_src + done, static_cast<size_t>(n));
            GGRM_CUDA(cudaMemcpyAsync(dst + done, s->bounce[which], static_cast<size_t>(n), cudaMemcpyHostToDevice, s->stream));
            GGRM_CUDA(cudaEventRecord(s->bounce_free[which], s->stream));
            done += n;
        }
    }
    GGRM_CUDA(cudaStreamSynchronize(s->stream));  // on return the slot holds the bytes: any handle may ggrm_push_device from it
    GGRM_STAGE_END(s)
}

extern "C" int ggrm_stage_slot(ggrm_stage_t* ss, int slot, const uint8_t** d_ptr) {
    Stage* s = reinterpret_cast<Stage*>(ss);
    GGRM_STAGE_BEGIN(s)
    if (slot < 0 || slot >= static_cast<int>(s->slots.size()) || d_ptr == nullptr) throw CudaError{GGRM_ERR_INVALID_ARGUMENT, "stage: slot out of range"};
    *d_ptr = s->slots[static_cast<size_t>(slot)];
    GGRM_STAGE_END(s)
}

// Every copy ggrm_push / ggrm_push_device has enqueued so far has read its source: a staging slot (or a caller-pinned buffer)
// handed to earlier pushes may be overwritten.
extern "C" int ggrm_wait_copies(ggrm_t* hh) {
    Handle* h = reinterpret_cast<Handle*>(hh);
    GGRM_API_BEGIN(h)
    materialize_direct_src(*h);  // a payload registered in place is copied now: the caller wants its memory back
    GGRM_CUDA(cudaStreamSynchronize(h->s_copy));
    for (auto& ev : h->raw_free) GGRM_CUDA(cudaEventSynchronize(ev));  // ... and a fused prep launch that reads one in place has run
    GGRM_API_END(h)
}
