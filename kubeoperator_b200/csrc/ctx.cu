// Life cycle, error reporting, buffers and pinning of libkorollup.so.
#include <stdarg.h>
#include <stdio.h>

#include "ko_common.cuh"

static thread_local char g_err[512] = "";

void ko_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int ko_buf_reserve(KoBuf* b, size_t bytes) {
    if (bytes <= b->cap) return KO_OK;
    if (b->p) {
        KO_CUDA(cudaFree(b->p));
        b->p = nullptr;
        b->cap = 0;
    }
    size_t want = ko_align_up(bytes + bytes / 8, 1 << 20);
    cudaError_t e = cudaMalloc(&b->p, want);
    if (e != cudaSuccess) {
        b->p = nullptr;
        ko_set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return KO_ENOMEM;
    }
    b->cap = want;
    return KO_OK;
}

int ko_pinned_reserve(ko_ctx* ctx, size_t bytes) {
    if (bytes <= ctx->pinned_out_cap) return KO_OK;
    if (ctx->pinned_out) {
        KO_CUDA(cudaFreeHost(ctx->pinned_out));
        ctx->pinned_out = nullptr;
        ctx->pinned_out_cap = 0;
    }
    size_t want = ko_align_up(bytes + bytes / 8, 1 << 16);
    cudaError_t e = cudaMallocHost(&ctx->pinned_out, want);
    if (e != cudaSuccess) {
        ctx->pinned_out = nullptr;
        ko_set_error("cudaMallocHost(%zu) failed: %s", want, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return KO_ENOMEM;
    }
    ctx->pinned_out_cap = want;
    return KO_OK;
}

extern "C" {

int ko_version(void) { return KO_VERSION_INT; }

const char* ko_last_error(void) { return g_err; }

int ko_init(const int* devices, int n_devices, ko_ctx** out) {
    KO_REQUIRE(out != nullptr, "ko_init: out is NULL");
    *out = nullptr;
    KO_REQUIRE(n_devices == 1 && devices != nullptr,
               "ko_init: one process drives one GPU (n_devices must be 1)");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        ko_set_error("ko_init: no CUDA device (%s); libkorollup has no CPU fallback",
                     e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        (void)cudaGetLastError();
        return KO_ECUDA;
    }
    KO_REQUIRE(devices[0] >= 0 && devices[0] < count, "ko_init: device ordinal out of range");
    KO_CUDA(cudaSetDevice(devices[0]));
    cudaDeviceProp prop;
    KO_CUDA(cudaGetDeviceProperties(&prop, devices[0]));
    if (prop.major != 10) {
        ko_set_error("ko_init: device %d is sm_%d%d; libkorollup is built for sm_100a (B200) only",
                     devices[0], prop.major, prop.minor);
        return KO_ECUDA;
    }
    ko_ctx* ctx = new ko_ctx();
    ctx->device = devices[0];
    ctx->num_sms = prop.multiProcessorCount;
    // streams and events; on any failure what exists so far is destroyed again and *out stays NULL
    auto create_all = [&]() -> cudaError_t {
        cudaError_t r;
        if ((r = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return r;
        if ((r = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return r;
        for (int i = 0; i < 2; ++i) {
            if ((r = cudaEventCreateWithFlags(&ctx->ev_copied[i], cudaEventDisableTiming)) != cudaSuccess) return r;
            if ((r = cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming)) != cudaSuccess) return r;
        }
        if ((r = cudaEventCreate(&ctx->ev_k0)) != cudaSuccess) return r;
        if ((r = cudaEventCreate(&ctx->ev_k1)) != cudaSuccess) return r;
        if ((r = cudaEventCreate(&ctx->ev_p0)) != cudaSuccess) return r;
        return cudaEventCreate(&ctx->ev_p1);
    };
    const cudaError_t ce = create_all();
    if (ce != cudaSuccess) {
        ko_set_error("ko_init: creating streams / events failed: %s", cudaGetErrorString(ce));
        for (int i = 0; i < 2; ++i) {
            if (ctx->ev_copied[i]) cudaEventDestroy(ctx->ev_copied[i]);
            if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
        }
        cudaEvent_t evs[] = {ctx->ev_k0, ctx->ev_k1, ctx->ev_p0, ctx->ev_p1};
        for (cudaEvent_t ev : evs)
            if (ev) cudaEventDestroy(ev);
        if (ctx->stream) cudaStreamDestroy(ctx->stream);
        if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
        delete ctx;
        (void)cudaGetLastError();
        return ce == cudaErrorMemoryAllocation ? KO_ENOMEM : KO_ECUDA;
    }
    ctx->alive = true;
    *out = ctx;
    return KO_OK;
}

void ko_destroy(ko_ctx* ctx) {
    if (!ctx) return;
    {
        std::lock_guard<std::mutex> lk(ctx->mu);
        if (!ctx->alive) return;
        ctx->alive = false;
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        cudaStreamSynchronize(ctx->copy_stream);
        KoBuf* bufs[] = {&ctx->scratch, &ctx->stage[0], &ctx->stage[1], &ctx->in_small, &ctx->out_small};
        for (KoBuf* b : bufs)
            if (b->p) cudaFree(b->p);
        if (ctx->pinned_out) cudaFreeHost(ctx->pinned_out);
        for (int i = 0; i < 2; ++i) {
            cudaEventDestroy(ctx->ev_copied[i]);
            cudaEventDestroy(ctx->ev_done[i]);
        }
        cudaEventDestroy(ctx->ev_k0);
        cudaEventDestroy(ctx->ev_k1);
        cudaEventDestroy(ctx->ev_p0);
        cudaEventDestroy(ctx->ev_p1);
        cudaStreamDestroy(ctx->stream);
        cudaStreamDestroy(ctx->copy_stream);
    }
    delete ctx;
}

int ko_get_counters(ko_ctx* ctx, ko_counters* out) {
    KO_REQUIRE(ctx && out, "ko_get_counters: NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    *out = ctx->counters;
    return KO_OK;
}

int ko_last_series_kernel_ms(ko_ctx* ctx, float* out_ms) {
    KO_REQUIRE(ctx && out_ms, "ko_last_series_kernel_ms: NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_REQUIRE(ctx->k_timed, "ko_last_series_kernel_ms: no series rollup has been launched yet");
    KO_CUDA(cudaEventSynchronize(ctx->ev_k1));
    KO_CUDA(cudaEventElapsedTime(out_ms, ctx->ev_k0, ctx->ev_k1));
    return KO_OK;
}

int ko_last_series_pipe_ms(ko_ctx* ctx, float* out_ms) {
    KO_REQUIRE(ctx && out_ms, "ko_last_series_pipe_ms: NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_REQUIRE(ctx->k_timed, "ko_last_series_pipe_ms: no series rollup has been launched yet");
    KO_CUDA(cudaEventSynchronize(ctx->ev_k1));
    KO_CUDA(cudaEventElapsedTime(out_ms, ctx->ev_p0, ctx->ev_p1));
    return KO_OK;
}

int ko_pin(void* ptr, int64_t bytes) {
    KO_REQUIRE(ptr && bytes > 0, "ko_pin: bad argument");
    KO_CUDA(cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterDefault));
    return KO_OK;
}

int ko_unpin(void* ptr) {
    KO_REQUIRE(ptr, "ko_unpin: NULL");
    KO_CUDA(cudaHostUnregister(ptr));
    return KO_OK;
}

}  // extern "C"
