// Context management of librpvg_b200.so.

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

extern "C" {

const char * rpvg_b200_version(void) {

    return "rpvg_b200 0.1 (sm_100a)";
}

int rpvg_b200_device_count(void) {

    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return count;
}

int rpvg_b200_create(int device, void * stream, rpvg_b200_ctx ** out_ctx) {

    if (!out_ctx) { return RPVG_B200_ERR_INVALID; }
    *out_ctx = nullptr;

    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);

    // No device, no library: there is no CPU fallback behind this ABI.
    if (err != cudaSuccess || count == 0) { return RPVG_B200_ERR_NO_DEVICE; }
    if (device < 0 || device >= count) { return RPVG_B200_ERR_INVALID; }

    if (cudaSetDevice(device) != cudaSuccess) { return RPVG_B200_ERR_CUDA; }

#ifdef RPVG_B200_EM_TIMELINE
    cudaDeviceSetLimit(cudaLimitPrintfFifoSize, 64u << 20);   // one line per CTA
#endif

    auto * ctx = new rpvg_b200_ctx();
    ctx->device = device;

    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        delete ctx;
        return RPVG_B200_ERR_CUDA;
    }

    ctx->num_sms = prop.multiProcessorCount;
    ctx->max_smem_optin = static_cast<int>(prop.sharedMemPerBlockOptin);

    if (stream) {

        ctx->stream = static_cast<cudaStream_t>(stream);
        ctx->own_stream = false;

    } else {

        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete ctx;
            return RPVG_B200_ERR_CUDA;
        }
        ctx->own_stream = true;
    }

    // A private memory pool that keeps freed batch memory: batches of similar size follow each other.
    {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;

        if (cudaMemPoolCreate(&ctx->pool, &props) != cudaSuccess) {
            if (ctx->own_stream) { cudaStreamDestroy(ctx->stream); }
            delete ctx;
            return RPVG_B200_ERR_CUDA;
        }

        uint64_t threshold = UINT64_MAX;
        cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &threshold);
    }

    // Side streams for the concurrent launches of rpvg_b200_batch_em, two per priority level below the lanes' and four at
    // the lowest (numerically lower = higher priority). The work distributor hands out the CTAs of one kernel per
    // priority level at a time: launches that are to overlap with a kernel that still has undispatched thread-block
    // clusters must sit on another level.
    int prio_low = 0, prio_high = 0;
    cudaDeviceGetStreamPriorityRange(&prio_low, &prio_high);
    if (std::getenv("RPVG_B200_DEBUG")) { fprintf(stderr, "[rpvg_b200] stream priorities %d (lowest) .. %d (highest)\n", prio_low, prio_high); }

    for (int k = 0; k < kNumSideStreams; ++k) {

        cudaStream_t side = nullptr;
        cudaEvent_t join = nullptr;

        if (cudaStreamCreateWithPriority(&side, cudaStreamNonBlocking, std::min(prio_low, prio_high + 1 + std::min(k / 2, 4))) != cudaSuccess ||
            cudaEventCreateWithFlags(&join, cudaEventDisableTiming) != cudaSuccess) {
            rpvg_b200_destroy(ctx);
            return RPVG_B200_ERR_CUDA;
        }

        ctx->side_streams.push_back(side);
        ctx->join_events.push_back(join);
    }

    for (int k = 0; k < 8; ++k) {

        cudaStream_t lane = nullptr;
        cudaEvent_t done = nullptr;

        if (cudaStreamCreateWithPriority(&lane, cudaStreamNonBlocking, prio_high) != cudaSuccess ||
            cudaEventCreateWithFlags(&done, cudaEventDisableTiming) != cudaSuccess) {
            rpvg_b200_destroy(ctx);
            return RPVG_B200_ERR_CUDA;
        }

        ctx->lane_streams.push_back(lane);
        ctx->lane_events.push_back(done);
    }

    if (cudaEventCreateWithFlags(&ctx->fork_event, cudaEventDisableTiming) != cudaSuccess || rpvgEmInit(ctx) != RPVG_B200_OK) {
        rpvg_b200_destroy(ctx);
        return RPVG_B200_ERR_CUDA;
    }

    *out_ctx = ctx;
    return RPVG_B200_OK;
}

void rpvg_b200_destroy(rpvg_b200_ctx * ctx) {

    if (!ctx) { return; }

    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto side: ctx->side_streams) {
        cudaStreamSynchronize(side);
        cudaStreamDestroy(side);
    }
    for (auto ev: ctx->join_events) { cudaEventDestroy(ev); }
    for (auto lane: ctx->lane_streams) {
        cudaStreamSynchronize(lane);
        cudaStreamDestroy(lane);
    }
    for (auto ev: ctx->lane_events) { cudaEventDestroy(ev); }
    if (ctx->fork_event) { cudaEventDestroy(ctx->fork_event); }
    if (ctx->own_stream) { cudaStreamDestroy(ctx->stream); }
    if (ctx->pool) { cudaMemPoolDestroy(ctx->pool); }
    for (auto buffer: ctx->pinned) {
        if (buffer) { cudaFreeHost(buffer); }
    }

    delete ctx;
}

void * rpvg_b200_pinned_scratch(rpvg_b200_ctx * ctx, uint32_t slot, uint64_t bytes) {

    if (!ctx || slot >= rpvg_b200_ctx::kPinnedSlots) {
        rpvgFail(ctx, RPVG_B200_ERR_INVALID, "rpvg_b200_pinned_scratch: bad slot");
        return nullptr;
    }

    if (bytes <= ctx->pinned_bytes[slot] && ctx->pinned[slot]) { return ctx->pinned[slot]; }

    cudaSetDevice(ctx->device);

    if (ctx->pinned[slot]) {
        cudaStreamSynchronize(ctx->stream);   // a copy from the old buffer may still be in flight
        cudaFreeHost(ctx->pinned[slot]);
        ctx->pinned[slot] = nullptr;
        ctx->pinned_bytes[slot] = 0;
    }

    // a quarter of headroom: batches of a run are of similar, not equal, size
    const uint64_t want = std::max<uint64_t>(bytes + bytes / 4, 1 << 16);
    void * buffer = nullptr;
    const cudaError_t err = cudaHostAlloc(&buffer, want, cudaHostAllocDefault);

    if (err != cudaSuccess) {
        rpvgFail(ctx, RPVG_B200_ERR_CUDA, std::string("cudaHostAlloc(") + std::to_string(want) + "): " + cudaGetErrorString(err));
        cudaGetLastError();
        return nullptr;
    }

    ctx->pinned[slot] = buffer;
    ctx->pinned_bytes[slot] = want;
    return buffer;
}

const char * rpvg_b200_last_error(const rpvg_b200_ctx * ctx) {

    return ctx ? ctx->error.c_str() : "no context (is a CUDA device present?)";
}

uint64_t rpvg_b200_launch_count(const rpvg_b200_ctx * ctx) {

    return ctx ? ctx->launches : 0;
}

}
