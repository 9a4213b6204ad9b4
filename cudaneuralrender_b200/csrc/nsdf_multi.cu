// =====================================================================================
// nsdf_multi.cu -- one process, several GPUs (include/nsdf.h, nsdf_multi_*).
//
// The reference is single-GPU.  Every pixel is independent (singleMarch touches only index `id`,
// src/volumeRender_kernel.cu:416-477), so a frame shards by 64x64 tiles and a turntable (src/main.cpp:470-478) by views,
// with no data-path collective:
//   * one view  : device i renders the tiles t with t % n == i and stores them straight into the presenting device's frame
//                 through a peer mapping (cudaDeviceEnablePeerAccess): finished tiles cross NVLink as 256-byte row
//                 segments from inside the tracing kernel (render_kernels.cuh flush_tile); one stream sync per device is
//                 the completion flag;
//   * n views   : view v goes to device v % n; each device renders its views in ONE persistent launch into local memory
//                 and copies them to the presenting device with cudaMemcpyPeerAsync on its own stream.
// All launches are issued from the calling host thread, device by device (they are asynchronous), then joined.
// =====================================================================================
#include <cstring>
#include <vector>

#include "nsdf_internal.h"

struct nsdf_multi {
    std::vector<int> devices;
    std::vector<nsdf_model*> models;         // one clone of the network per device
    std::vector<cudaStream_t> streams;
    std::vector<uint32_t*> d_matcap;         // per device
    std::vector<size_t> matcap_bytes;
    std::vector<uint32_t*> d_local;          // per device: frames of its views (view mode)
    std::vector<size_t> local_bytes;
    std::vector<nsdf_render_stats*> d_stats; // per device
};

namespace {

struct DeviceGuard {
    int prev = 0;
    DeviceGuard() { cudaGetDevice(&prev); }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

#define NSDF_MTRY(expr)                          \
    do {                                         \
        cudaError_t e__ = (expr);                \
        if (e__ != cudaSuccess) return (int)e__; \
    } while (0)

int grow(void** p, size_t* have, size_t need) {
    if (need <= *have) return NSDF_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    NSDF_MTRY(cudaMalloc(p, need));
    *have = need;
    return NSDF_OK;
}

}  // namespace

extern "C" {

int nsdf_multi_create(const nsdf_model* m, const int* devices, int n_devices, nsdf_multi** out) {
    if (!m || !devices || n_devices < 1 || n_devices > 64 || !out) return NSDF_EINVAL;
    *out = nullptr;
    int count = 0;
    NSDF_MTRY(cudaGetDeviceCount(&count));
    for (int i = 0; i < n_devices; ++i) {
        if (devices[i] < 0 || devices[i] >= count) return NSDF_EINVAL;
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i]) return NSDF_EINVAL;
    }
    DeviceGuard guard;
    nsdf_multi* mm = new (std::nothrow) nsdf_multi();
    if (!mm) return NSDF_ENOMEM;
    std::vector<int> dims((size_t)m->n_layers + 1);
    dims[0] = m->in[0];
    for (int l = 0; l < m->n_layers; ++l) dims[(size_t)l + 1] = m->out[l];
    int rc = NSDF_OK;
    for (int i = 0; i < n_devices && rc == NSDF_OK; ++i) {
        cudaError_t e = cudaSetDevice(devices[i]);
        if (e != cudaSuccess) { rc = (int)e; break; }
        nsdf_model* clone = nullptr;
        rc = nsdf_model_create_packed(m->h_params.data(), dims.data(), m->n_layers, &clone);
        if (rc) break;
        cudaStream_t st = nullptr;
        e = cudaStreamCreate(&st);   // blocking: ordered after the legacy default stream, like the synchronous call this serves
        if (e != cudaSuccess) { nsdf_model_destroy(clone); rc = (int)e; break; }
        nsdf_render_stats* ds = nullptr;
        e = cudaMalloc(&ds, sizeof(nsdf_render_stats));
        if (e != cudaSuccess) { nsdf_model_destroy(clone); cudaStreamDestroy(st); rc = (int)e; break; }
        mm->devices.push_back(devices[i]);
        mm->models.push_back(clone);
        mm->streams.push_back(st);
        mm->d_matcap.push_back(nullptr);
        mm->matcap_bytes.push_back(0);
        mm->d_local.push_back(nullptr);
        mm->local_bytes.push_back(0);
        mm->d_stats.push_back(ds);
        if (i > 0) {   // peers store into / copy to the presenting device (devices[0])
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[i], devices[0]);
            if (can) {
                e = cudaDeviceEnablePeerAccess(devices[0], 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = (int)e;
                cudaGetLastError();
            } else {
                rc = NSDF_EUNSUPPORTED;
            }
        }
    }
    if (rc) {
        nsdf_multi_destroy(mm);
        return rc;
    }
    *out = mm;
    return NSDF_OK;
}

void nsdf_multi_destroy(nsdf_multi* mm) {
    if (!mm) return;
    DeviceGuard guard;
    for (size_t i = 0; i < mm->devices.size(); ++i) {
        cudaSetDevice(mm->devices[i]);
        if (mm->streams[i]) { cudaStreamSynchronize(mm->streams[i]); cudaStreamDestroy(mm->streams[i]); }
        if (mm->d_matcap[i]) cudaFree(mm->d_matcap[i]);
        if (mm->d_local[i]) cudaFree(mm->d_local[i]);
        if (mm->d_stats[i]) cudaFree(mm->d_stats[i]);
        nsdf_model_destroy(mm->models[i]);
    }
    delete mm;
}

int nsdf_multi_device_count(const nsdf_multi* mm) { return mm ? (int)mm->devices.size() : 0; }

int nsdf_multi_render(nsdf_multi* mm, const nsdf_camera* cams, int n_views, const uint32_t* h_matcap, int mw, int mh,
                      const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* h_stats) {
    if (!mm || !cams || n_views < 1 || !d_out || W <= 0 || H <= 0) return NSDF_EINVAL;
    nsdf_render_opts o;
    if (opts) o = *opts;
    else nsdf_render_opts_default(&o);
    if (o.tile_world != 1 || o.tile_rank != 0) return NSDF_EINVAL;   // the partition is this call's business
    const int n = (int)mm->devices.size();
    const bool have_mat = h_matcap && mw > 0 && mh > 0;
    const size_t frame_px = (size_t)W * H;
    DeviceGuard guard;
    // the call is synchronous: whatever the caller queued on the presenting device for d_out (a memset, a clear) comes first
    NSDF_MTRY(cudaSetDevice(mm->devices[0]));
    NSDF_MTRY(cudaDeviceSynchronize());
    // ---- issue: matcap upload + launch(es) on every device's own stream
    for (int i = 0; i < n; ++i) {
        NSDF_MTRY(cudaSetDevice(mm->devices[i]));
        cudaStream_t st = mm->streams[i];
        if (have_mat) {
            int rc = grow((void**)&mm->d_matcap[i], &mm->matcap_bytes[i], (size_t)mw * mh * sizeof(uint32_t));
            if (rc) return rc;
            NSDF_MTRY(cudaMemcpyAsync(mm->d_matcap[i], h_matcap, (size_t)mw * mh * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        }
        NSDF_MTRY(cudaMemsetAsync(mm->d_stats[i], 0, sizeof(nsdf_render_stats), st));
        const uint32_t* mat = have_mat ? mm->d_matcap[i] : nullptr;
        if (n_views == 1) {
            nsdf_render_opts oi = o;
            oi.tile_rank = i;
            oi.tile_world = n;
            const int rc = nsdf_render(mm->models[i], cams, mat, mw, mh, &oi, d_out, W, H, mm->d_stats[i], st);
            if (rc) return rc;
        } else {
            std::vector<nsdf_camera> mine;
            for (int v = i; v < n_views; v += n) mine.push_back(cams[v]);
            if (mine.empty()) continue;
            uint32_t* dst = d_out;
            if (i > 0 || n > 1) {   // render into local memory, then copy each view to its slot on the presenting device
                int rc = grow((void**)&mm->d_local[i], &mm->local_bytes[i], mine.size() * frame_px * sizeof(uint32_t));
                if (rc) return rc;
                dst = mm->d_local[i];
            }
            for (size_t s = 0; s < mine.size(); s += 128) {
                const int cnt = (int)(mine.size() - s < 128 ? mine.size() - s : 128);
                const int rc = nsdf_render_views(mm->models[i], mine.data() + s, cnt, mat, mw, mh, &o, dst + s * frame_px, W, H,
                                                 mm->d_stats[i], st);
                if (rc) return rc;
            }
            if (dst != d_out) {
                for (size_t k = 0; k < mine.size(); ++k) {
                    const size_t v = (size_t)i + k * (size_t)n;
                    NSDF_MTRY(cudaMemcpyPeerAsync(d_out + v * frame_px, mm->devices[0], dst + k * frame_px, mm->devices[i],
                                                  frame_px * sizeof(uint32_t), st));
                }
            }
        }
    }
    // ---- join
    nsdf_render_stats total;
    std::memset(&total, 0, sizeof(total));
    for (int i = 0; i < n; ++i) {
        NSDF_MTRY(cudaSetDevice(mm->devices[i]));
        NSDF_MTRY(cudaStreamSynchronize(mm->streams[i]));
        if (h_stats) {
            nsdf_render_stats s;
            NSDF_MTRY(cudaMemcpy(&s, mm->d_stats[i], sizeof(s), cudaMemcpyDeviceToHost));
            total.queries += s.queries; total.hit_pixels += s.hit_pixels; total.rays += s.rays;
            total.evicted += s.evicted; total.unfinished += s.unfinished; total.pixels += s.pixels;
        }
    }
    if (h_stats) *h_stats = total;
    return NSDF_OK;
}

}  // extern "C"
