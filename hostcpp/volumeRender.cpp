// Host side of the reference's render entry points (include/volumeRender.hh) over the C ABI.
// State that the reference keeps in __constant__ symbols (src/volumeRender_kernel.cu:31-35) lives here.
#include <cstring>
#include "volumeRender.hh"

namespace {
struct RenderState {
    nsdf_camera cam;
    nsdf_render_opts opts;
    nsdf_render_stats last;
    nsdf_render_stats* d_stats = nullptr;
    bool init = false;
} g;

void ensure_init() {
    if (g.init) return;
    std::memset(&g.cam, 0, sizeof(g.cam));
    nsdf_render_opts_default(&g.opts);
    std::memset(&g.last, 0, sizeof(g.last));
    nsdf_camera_from_view(0.f, 0.f, 0.f, 0.f, -2.f, 0, &g.cam);
    g.init = true;
}
}  // namespace

extern "C" void render_kernel(dim3, dim3, uint* d_output, uint imageW, uint imageH, uint numInputs, NeuralNetwork& nn,
                              Image matcap) {
    ensure_init();
    (void)numInputs;
    nsdf_model* m = nn.handle();
    if (!m) {
        std::fprintf(stderr, "render_kernel: the network has no dense layers\n");
        std::exit(EXIT_FAILURE);
    }
    if (!g.d_stats) nsdfCheck(nsdf_dev_malloc(reinterpret_cast<void**>(&g.d_stats), sizeof(nsdf_render_stats)));
    nsdfCheck(nsdf_dev_memset(g.d_stats, 0, sizeof(nsdf_render_stats)));
    const bool useMatcap = g.opts.color_type == NSDF_COLOR_MATCAP;
    nsdfCheck(nsdf_render(m, &g.cam, useMatcap ? matcap.deviceData.get() : nullptr, (int)matcap.shape.x, (int)matcap.shape.y,
                          &g.opts, d_output, (int)imageW, (int)imageH, g.d_stats, nullptr));
    nsdfCheck(nsdf_memcpy_d2h(&g.last, g.d_stats, sizeof(nsdf_render_stats)));   // also the frame's sync point
}

extern "C" void copyViewMatrices(float* invViewMatrix, size_t sizeofInvModelViewMat, float* normalMatrix,
                                 size_t sizeofNormalMatrix, int frameNumber, int) {
    ensure_init();
    std::memcpy(g.cam.inv_view, invViewMatrix, sizeofInvModelViewMat < sizeof(g.cam.inv_view) ? sizeofInvModelViewMat : sizeof(g.cam.inv_view));
    std::memcpy(g.cam.normal, normalMatrix, sizeofNormalMatrix < sizeof(g.cam.normal) ? sizeofNormalMatrix : sizeof(g.cam.normal));
    g.cam.frame = frameNumber;
}

extern "C" void copyStaticSettings(int colorType, int) {
    ensure_init();
    g.opts.color_type = colorType;
}

extern "C" void setRenderOptions(int scene, int maxSteps, int flags) {
    ensure_init();
    g.opts.scene = scene;
    g.opts.max_steps = maxSteps;
    g.opts.flags = flags;
}

extern "C" void setMathMode(int math) {
    ensure_init();
    g.opts.math = math;
}

extern "C" void setSceneProgram(const nsdf_scene_op* ops, int nOps) {
    ensure_init();
    static nsdf_scene_op program[NSDF_SCENE_MAX_OPS];
    if (!ops || nOps < 1 || nOps > NSDF_SCENE_MAX_OPS) {
        std::fprintf(stderr, "setSceneProgram: 1..%d ops expected\n", NSDF_SCENE_MAX_OPS);
        std::exit(EXIT_FAILURE);
    }
    std::memcpy(program, ops, sizeof(nsdf_scene_op) * (size_t)nOps);
    g.opts.scene = NSDF_SCENE_PROGRAM;
    g.opts.program = program;
    g.opts.n_ops = nOps;
}

extern "C" void getRenderOptions(nsdf_render_opts* out) {
    ensure_init();
    *out = g.opts;
}

extern "C" void getLastRenderStats(nsdf_render_stats* out) {
    ensure_init();
    *out = g.last;
}

extern "C" void updateViewMatrices(float rxDeg, float ryDeg, float tx, float ty, float tz, int frameNumber) {
    ensure_init();
    nsdf_camera_from_view(rxDeg, ryDeg, tx, ty, tz, frameNumber, &g.cam);
}
