// Host side of mlp_tc.cu: weight-image builder, launcher, rung dispatch.  Included by mlp_tc.cu only.
namespace nsdf {
namespace {

#ifdef NSDF_DEV_SCHEDULES
int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}
#else
constexpr int env_int(const char*, int dflt) { return dflt; }   // product build: no environment reads, measured defaults
#endif

// 16-bit encodings (round to nearest even) and their values
uint16_t enc16(float v, int mode) {
    uint16_t bits;
    if (mode == MODE_BF16) {
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        std::memcpy(&bits, &h, 2);
    } else {
        const __half h = __float2half_rn(v);
        std::memcpy(&bits, &h, 2);
    }
    return bits;
}
float dec16(uint16_t bits, int mode) {
    if (mode == MODE_BF16) {
        return __bfloat162float(__nv_bfloat16(__nv_bfloat16_raw{bits}));
    }
    return __half2float(__half(__half_raw{bits}));
}

struct ImageWriter {
    std::vector<uint8_t>& img;
    int mode;
    void put(uint32_t base, int n, int k, int rows, uint16_t bits) {
        std::memcpy(img.data() + base + tc::b_elem_offset(n, k, rows), &bits, 2);
    }
    // writes hi at (n,k) of `base` and returns the residual v - hi
    float put_hi(uint32_t base, int n, int k, int rows, float v) {
        const uint16_t b = enc16(v, mode);
        put(base, n, k, rows, b);
        return v - dec16(b, mode);
    }
};

// W of layer l is [out][in] row-major in h_params (DenseLayer::W, src/layers/denseLayer.cu:217-227).
std::vector<uint8_t> build_image(const nsdf_model* m, int mode) {
    const int n_sq = m->n_sq;
    const ImageLayout L = tc::image_layout(n_sq, mode);
    std::vector<uint8_t> img(L.bytes, 0);
    ImageWriter w{img, mode};
    const float* P = m->h_params.data();
    {   // layer 0: k0-2 = wh, k3-5 = wh (multiplies x_lo), k6-8 = wl (multiplies x_hi), k9 = bh, k10 = bl
        const int in0 = m->in[0];
        for (int n = 0; n < 32; ++n) {
            for (int c = 0; c < 3 && c < in0; ++c) {
                const float v = P[m->w_off[0] + (size_t)n * in0 + c];
                const float r = w.put_hi(L.b0, n, c, 32, v);
                w.put_hi(L.b0, n, 3 + c, 32, v);
                w.put_hi(L.b0, n, 6 + c, 32, r);
            }
            const float rb = w.put_hi(L.b0, n, 9, 32, P[m->b_off[0] + n]);
            w.put_hi(L.b0, n, 10, 32, rb);
        }
    }
    for (int l = 0; l < n_sq; ++l) {
        const float* W = P + m->w_off[1 + l];
        const uint32_t bh = L.hid_hi + (uint32_t)l * tc::IMG_HID_HI_BYTES;
        const uint32_t bl = L.hid_lo + (uint32_t)l * tc::IMG_HID_LO_BYTES;
        for (int n = 0; n < 32; ++n) {
            for (int k = 0; k < 32; ++k) {
                const float r = w.put_hi(bh, n, k, 32, W[n * 32 + k]);
                if (mode == MODE_FP16X3) w.put_hi(bl, n, k, 32, r);
            }
            const float rb = w.put_hi(bh, n, 32, 32, P[m->b_off[1 + l] + n]);
            w.put_hi(bh, n, 33, 32, rb);
        }
    }
    {   // last layer: one output neuron in row 0 of a 16-row operand
        const int last = m->n_layers - 1;
        for (int k = 0; k < 32; ++k) {
            const float r = w.put_hi(L.last_hi, 0, k, 16, P[m->w_off[last] + k]);
            if (mode == MODE_FP16X3) w.put_hi(L.last_lo, 0, k, 16, r);
        }
        const float rb = w.put_hi(L.last_hi, 0, 32, 16, P[m->b_off[last]]);
        w.put_hi(L.last_hi, 0, 33, 16, rb);
    }
    return img;
}

int ensure_images(nsdf_model* m) {
    if (m->d_tc) return NSDF_OK;
    TcDeviceImages* t = new TcDeviceImages();
    size_t total = 0;
    std::vector<std::vector<uint8_t>> imgs;
    for (int mode = 0; mode < N_MODES; ++mode) {
        imgs.push_back(build_image(m, mode));
        t->bytes[mode] = (uint32_t)imgs.back().size();
        total += imgs.back().size();
    }
    uint8_t* d = nullptr;
    cudaError_t e = cudaMalloc(&d, total);
    if (e != cudaSuccess) { delete t; return (int)e; }
    size_t off = 0;
    for (int mode = 0; mode < N_MODES; ++mode) {
        t->d[mode] = d + off;
        e = cudaMemcpy(d + off, imgs[mode].data(), imgs[mode].size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { cudaFree(d); delete t; return (int)e; }
        off += imgs[mode].size();
    }
    m->d_tc = t;
    m->tc_bytes = total;
    return NSDF_OK;
}

long long* g_tc_trace = nullptr;   // set by nsdf_debug_tc_trace (development)

template <int MODE, int HALVES, bool ACC16 = false>
int launch(nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
           cudaStream_t stream) {
    TcParams p;
    p.image = t->d[MODE];
    p.image_bytes = t->bytes[MODE];
    p.n_sq = m->n_sq;
    p.x = d_x;
    p.y = d_y;
    p.n = n;
    p.lp = lp;
    p.trace = g_tc_trace;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(mlp_tc_kernel<MODE, HALVES, ACC16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    // Resident CTAs per SM = 512 TMEM columns / columns per CTA.  Shared memory is padded so that no
    // more CTAs than that are ever co-resident (tcgen05.alloc would otherwise spin).
    const int max_per_sm = 512 / (int)tc::Cols<MODE>::TOTAL;
    int per_sm = env_int("NSDF_TC_CTAS_PER_SM", max_per_sm);
    if (per_sm > max_per_sm) per_sm = max_per_sm;
    if (per_sm < 1) per_sm = 1;
    size_t smem = p.image_bytes;
    const size_t min_smem = (size_t)(227 * 1024) / (size_t)(per_sm + 1) + 1;
    if (smem < min_smem) smem = min_smem;
    if (smem > 120 * 1024) return NSDF_EUNSUPPORTED;
    const long long n_tiles = (n + 127) / 128;
    const long long cap = (long long)sms * per_sm;
    const int grid = (int)(n_tiles < cap ? n_tiles : cap);
    mlp_tc_kernel<MODE, HALVES, ACC16><<<grid, TC_THREADS, smem, stream>>>(p);
    return (int)cudaGetLastError();
}

// ---- v3: warp-specialised persistent kernel (tc_mlp3.cuh) ------------------------------------------------
template <int MODE, int GW, int T, int NM>
int launch3(nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
            cudaStream_t stream) {
    tc::Tc3Params p;
    p.image = t->d[MODE];
    p.image_bytes = t->bytes[MODE];
    p.n_sq = m->n_sq;
    p.x = d_x;
    p.y = d_y;
    p.n = n;
    p.lp = lp;
    p.trace = g_tc_trace;
    const size_t smem = 120 * 1024;   // > half of the SM's shared memory: exactly one CTA (and its 512 TMEM columns) per SM
    if (p.image_bytes > smem) return NSDF_EUNSUPPORTED;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc::mlp_tc3_kernel<MODE, GW, T, NM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    const long long n_tiles = (n + 127) / 128;
    const long long want = (n_tiles + GW * T - 1) / (GW * T);
    const int grid = (int)(want < sms ? want : sms);
    tc::mlp_tc3_kernel<MODE, GW, T, NM><<<grid, (GW * 4 + NM) * 32, smem, stream>>>(p);
    return (int)cudaGetLastError();
}

template <int MODE>
int launch3_cfg(int cfg, nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
                cudaStream_t stream) {
    constexpr bool X3 = MODE == MODE_FP16X3;   // 64-column slots: at most 7 tiles in flight
    switch (cfg) {
        case 0: return launch3<MODE, 7, 1, 1>(m, t, d_x, d_y, n, lp, stream);
        case 1: return launch3<MODE, X3 ? 3 : 4, 2, 1>(m, t, d_x, d_y, n, lp, stream);
        case 2: return launch3<MODE, 3, X3 ? 2 : 3, 1>(m, t, d_x, d_y, n, lp, stream);
        case 3: return launch3<MODE, X3 ? 3 : 5, 2, 2>(m, t, d_x, d_y, n, lp, stream);
        case 4: return launch3<MODE, X3 ? 3 : 4, 2, 2>(m, t, d_x, d_y, n, lp, stream);
        case 5: return launch3<MODE, 7, 1, 2>(m, t, d_x, d_y, n, lp, stream);
        case 6: return launch3<MODE, X3 ? 3 : 5, 2, 1>(m, t, d_x, d_y, n, lp, stream);
        case 7: return launch3<MODE, 7, 1, 4>(m, t, d_x, d_y, n, lp, stream);
        default: return NSDF_EINVAL;
    }
}

// ---- v4: grouped CTAs, bias pre-loaded into the accumulator, UMMAs issued in parallel (tc_mlp4.cuh) ---------
template <int MODE, int G, int CPS, bool SP = false>
int launch4(nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
            cudaStream_t stream) {
    tc::Tc4Params p;
    p.image = t->d[MODE];
    p.image_bytes = t->bytes[MODE];
    p.blob = m->d_blob;
    p.n_sq = m->n_sq;
    p.x = d_x;
    p.y = d_y;
    p.n = n;
    p.lp = lp;
    size_t smem = p.image_bytes;
    const size_t min_smem = (size_t)(227 * 1024) / (size_t)(CPS + 1) + 1;   // never more than CPS CTAs (and their TMEM) per SM
    if (smem < min_smem) smem = min_smem;
    if (smem > 200 * 1024) return NSDF_EUNSUPPORTED;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc::mlp_tc4_kernel<MODE, G, CPS, SP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    const long long n_tiles = (n + 127) / 128;
    const long long want = (n_tiles + G - 1) / G;
    const long long cap = (long long)sms * CPS;
    const int grid = (int)(want < cap ? want : cap);
    tc::mlp_tc4_kernel<MODE, G, CPS, SP><<<grid, G * 128, smem, stream>>>(p);
    return (int)cudaGetLastError();
}

template <int MODE, int G, int CPS>
int launch5(nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp, cudaStream_t stream) {
    tc::Tc4Params p;
    p.image = t->d[MODE];
    p.image_bytes = t->bytes[MODE];
    p.blob = m->d_blob;
    p.n_sq = m->n_sq;
    p.x = d_x;
    p.y = d_y;
    p.n = n;
    p.lp = lp;
    size_t smem = p.image_bytes;
    const size_t min_smem = (size_t)(227 * 1024) / (size_t)(CPS + 1) + 1;   // never more than CPS CTAs (and their TMEM) per SM
    if (smem < min_smem) smem = min_smem;
    if (smem > 200 * 1024) return NSDF_EUNSUPPORTED;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc::mlp_tc5_kernel<MODE, G, CPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    const long long n_tiles = (n + 127) / 128;
    const long long want = (n_tiles + G - 1) / G;
    const long long cap = (long long)sms * CPS;
    const int grid = (int)(want < cap ? want : cap);
    tc::mlp_tc5_kernel<MODE, G, CPS><<<grid, G * 128, smem, stream>>>(p);
    return (int)cudaGetLastError();
}

#ifdef NSDF_DEV_SCHEDULES      // v6 (two tile slots per warpgroup) measured slower than v5: development builds only
template <int MODE, int G, int CPS>
int launch6(nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp, cudaStream_t stream) {
    tc::Tc4Params p;
    p.image = t->d[MODE];
    p.image_bytes = t->bytes[MODE];
    p.blob = m->d_blob;
    p.n_sq = m->n_sq;
    p.x = d_x;
    p.y = d_y;
    p.n = n;
    p.lp = lp;
    size_t smem = p.image_bytes;
    const size_t min_smem = (size_t)(227 * 1024) / (size_t)(CPS + 1) + 1;   // never more than CPS CTAs (and their TMEM) per SM
    if (smem < min_smem) smem = min_smem;
    if (smem > 200 * 1024) return NSDF_EUNSUPPORTED;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc::mlp_tc6_kernel<MODE, G, CPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    const long long n_tiles = (n + 127) / 128;
    const long long want = (n_tiles + 2 * G - 1) / (2 * G);
    const long long cap = (long long)sms * CPS;
    const int grid = (int)(want < cap ? want : cap);
    tc::mlp_tc6_kernel<MODE, G, CPS><<<grid, G * 128, smem, stream>>>(p);
    return (int)cudaGetLastError();
}

template <int MODE>
int launch6_cfg(int cfg, nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
                cudaStream_t stream) {
    switch (cfg) {
        case 0: return launch6<MODE, 5, 1>(m, t, d_x, d_y, n, lp, stream);   // ten tiles in flight per SM, 640 threads
        case 1: return launch6<MODE, 2, 2>(m, t, d_x, d_y, n, lp, stream);   // eight, 512 threads
        case 2: return launch6<MODE, 1, 4>(m, t, d_x, d_y, n, lp, stream);   // eight, 512 threads in four CTAs
        case 3: return launch6<MODE, 4, 1>(m, t, d_x, d_y, n, lp, stream);   // eight, one CTA
        default: return NSDF_EINVAL;
    }
}

#endif

template <int MODE>
int launch5_cfg(int cfg, nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
                cudaStream_t stream) {
    switch (cfg) {
        case 0: return launch5<MODE, 5, 2>(m, t, d_x, d_y, n, lp, stream);   // ten tiles in flight per SM
        case 1: return launch5<MODE, 4, 2>(m, t, d_x, d_y, n, lp, stream);   // eight
        case 2: return launch5<MODE, 2, 4>(m, t, d_x, d_y, n, lp, stream);   // eight, in four CTAs
        case 3: return launch5<MODE, 1, 8>(m, t, d_x, d_y, n, lp, stream);   // eight, one per CTA
#ifdef NSDF_DEV_SCHEDULES      // rate against tiles in flight (profiles/r02_tc5_diag.txt)
        case 4: return launch5<MODE, 2, 3>(m, t, d_x, d_y, n, lp, stream);
        case 5: return launch5<MODE, 2, 2>(m, t, d_x, d_y, n, lp, stream);
        case 6: return launch5<MODE, 2, 1>(m, t, d_x, d_y, n, lp, stream);
        case 7: return launch5<MODE, 1, 4>(m, t, d_x, d_y, n, lp, stream);
#endif
        default: return NSDF_EINVAL;
    }
}

template <int MODE>
int launch4_cfg(int cfg, nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
                cudaStream_t stream) {
    constexpr bool X3 = MODE == MODE_FP16X3;   // 64-column slots
    switch (cfg) {
        case 0: return launch4<MODE, X3 ? 4 : 5, 2>(m, t, d_x, d_y, n, lp, stream);
        case 1: return launch4<MODE, 7, 1>(m, t, d_x, d_y, n, lp, stream);
        case 2: return launch4<MODE, 4, 2>(m, t, d_x, d_y, n, lp, stream);
        case 3: return launch4<MODE, 2, 4>(m, t, d_x, d_y, n, lp, stream);
        case 4: return launch4<MODE, 3, 2>(m, t, d_x, d_y, n, lp, stream);
        case 5: return launch4<MODE, 7, 1, true>(m, t, d_x, d_y, n, lp, stream);
        case 6: return launch4<MODE, X3 ? 4 : 5, 2, true>(m, t, d_x, d_y, n, lp, stream);
        case 7: return launch4<MODE, X3 ? 2 : 2, X3 ? 3 : 4, true>(m, t, d_x, d_y, n, lp, stream);
        default: return NSDF_EINVAL;
    }
}

}  // namespace

void tc_set_trace(long long* d_buf) { g_tc_trace = d_buf; }

const uint8_t* tc_image(nsdf_model* m, int mode, uint32_t* bytes) {
    if (ensure_images(m) != NSDF_OK) return nullptr;
    const TcDeviceImages* t = static_cast<const TcDeviceImages*>(m->d_tc);
    if (bytes) *bytes = t->bytes[mode];
    return t->d[mode];
}

int tc_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res, float lo,
               float hi, int64_t first, cudaStream_t stream) {
    if (n == 0) return NSDF_OK;
    if (m->n_sq < 1) return NSDF_EUNSUPPORTED;
    int rc = ensure_images(m);
    if (rc) return rc;
    const TcDeviceImages* t = static_cast<const TcDeviceImages*>(m->d_tc);
    LatticeParams lp{lattice ? 1 : 0, res, lo, hi, (long long)first};
    if (precision == NSDF_PREC_TC_FP16_ACC16) return launch<MODE_FP16, 1, true>(m, t, d_x, d_y, n, lp, stream);
    // Defaults from the measurements (profiles/r01_tc_schedule_notes.md, r02_tc5_diag.txt): the fp16/bf16 rungs are fastest on the v5
    // schedule with 2 tile groups x 4 CTAs per SM (41.6 Gq/s; v2 35.8, bit-equal); the fp16x3 rung (6 UMMAs per layer) on v4 with
    // 7 groups in one CTA per SM.
    const int kernel = env_int("NSDF_TC_KERNEL", precision == NSDF_PREC_TC_FP16X3 ? 4 : 5);
    if (kernel == 4) {
        const int cfg = env_int("NSDF_TC4_CFG", 5);   // 7 groups x 1 CTA, one polling warp per group (17.5 Gq/s; every warp polling: 17.0)
        switch (precision) {
            case NSDF_PREC_TC_FP16: return launch4_cfg<MODE_FP16>(cfg, m, t, d_x, d_y, n, lp, stream);
            case NSDF_PREC_TC_BF16: return launch4_cfg<MODE_BF16>(cfg, m, t, d_x, d_y, n, lp, stream);
            case NSDF_PREC_TC_FP16X3: return launch4_cfg<MODE_FP16X3>(cfg, m, t, d_x, d_y, n, lp, stream);
            default: return NSDF_EINVAL;
        }
    }
#ifdef NSDF_DEV_SCHEDULES
    if (kernel == 6 && precision != NSDF_PREC_TC_FP16X3) {
        const int cfg = env_int("NSDF_TC6_CFG", 0);
        return precision == NSDF_PREC_TC_FP16 ? launch6_cfg<MODE_FP16>(cfg, m, t, d_x, d_y, n, lp, stream)
                                              : launch6_cfg<MODE_BF16>(cfg, m, t, d_x, d_y, n, lp, stream);
    }
#endif
    if (kernel == 5 && precision != NSDF_PREC_TC_FP16X3) {
        const int cfg = env_int("NSDF_TC5_CFG", 2);
        return precision == NSDF_PREC_TC_FP16 ? launch5_cfg<MODE_FP16>(cfg, m, t, d_x, d_y, n, lp, stream)
                                              : launch5_cfg<MODE_BF16>(cfg, m, t, d_x, d_y, n, lp, stream);
    }
#ifdef NSDF_DEV_SCHEDULES      // the v3 schedule measured slower than v2 in every configuration: development builds only
    if (kernel == 3) {
        const int cfg = env_int("NSDF_TC3_CFG", 0);
        switch (precision) {
            case NSDF_PREC_TC_FP16: return launch3_cfg<MODE_FP16>(cfg, m, t, d_x, d_y, n, lp, stream);
            case NSDF_PREC_TC_BF16: return launch3_cfg<MODE_BF16>(cfg, m, t, d_x, d_y, n, lp, stream);
            case NSDF_PREC_TC_FP16X3: return launch3_cfg<MODE_FP16X3>(cfg, m, t, d_x, d_y, n, lp, stream);
            default: return NSDF_EINVAL;
        }
    }
#endif
    const int halves = env_int("NSDF_TC_EPI_HALVES", 1);
    switch (precision) {
        case NSDF_PREC_TC_FP16:
            return halves == 2 ? launch<MODE_FP16, 2>(m, t, d_x, d_y, n, lp, stream) : launch<MODE_FP16, 1>(m, t, d_x, d_y, n, lp, stream);
        case NSDF_PREC_TC_BF16:
            return halves == 2 ? launch<MODE_BF16, 2>(m, t, d_x, d_y, n, lp, stream) : launch<MODE_BF16, 1>(m, t, d_x, d_y, n, lp, stream);
        case NSDF_PREC_TC_FP16X3:
            return halves == 2 ? launch<MODE_FP16X3, 2>(m, t, d_x, d_y, n, lp, stream) : launch<MODE_FP16X3, 1>(m, t, d_x, d_y, n, lp, stream);
        default: return NSDF_EINVAL;
    }
}

void tc_release(nsdf_model* m) {
    if (m->d_tc) {
        TcDeviceImages* t = static_cast<TcDeviceImages*>(m->d_tc);
        if (t->d[0]) cudaFree(t->d[0]);
        delete t;
    }
    m->d_tc = nullptr;
    m->tc_bytes = 0;
}

}  // namespace nsdf
