// neuralSDFRenderer -- headless drop-in for the reference's renderer executable (src/main.cpp).
//
// Same flags as src/main.cpp:536-631 (-i -o -H -W -M -rx -ry -rz -z --single --spin --animation -h) and the
// same stdout line on the first frame (:434-438).  There is no window: without --spin one frame is rendered
// (the reference's --single path, generateSingleImage :404-468); --spin renders 360 frames with
// viewRotation.y = i and frameNumber = i (doABarrelRoll :470-478).  Output: <prefix><basename>.png with the
// reference's 180-degree turn (Image::savePNG), and the same frame as PPM with --ppm.
// Extra flags: --ppm, --scene manysphere|tanh|round|displace|sphere|cylindercut|spherecut (scene programs; default manysphere = HEAD), --math reference|deterministic, --max-steps N, --precision N (NSDF_PREC_*), --views N
// (turntable of N views instead of 360), --gpu K (device ordinal), --gpus N (one process drives N devices: a single frame is
// partitioned by 64x64 tiles, a turntable by views; nsdf_multi_*), --batch B (views per persistent launch and device,
// default 16), --stats.
// A turntable is rendered in batches (nsdf_render_views: the serial tail of one view hides behind the bulk of the others)
// while host threads encode and write the previous batch's PNG files, so the GPU never waits for the encoder.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <iostream>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "volumeRender.hh"

namespace {
// Frames handed to host threads: Image::savePNG (src/neuralUtils/image.cu:67-110) + the optional PPM, off the render loop.
class FrameWriter {
public:
    explicit FrameWriter(int n_threads) {
        for (int i = 0; i < n_threads; ++i) workers_.emplace_back([this] { run(); });
    }
    ~FrameWriter() { finish(); }
    void push(std::string name, std::shared_ptr<Image> img, bool ppm) {
        {
            std::lock_guard<std::mutex> lk(m_);
            q_.push_back(Job{std::move(name), std::move(img), ppm});
        }
        cv_.notify_one();
    }
    bool finish() {
        {
            std::lock_guard<std::mutex> lk(m_);
            if (done_) return ok_;
            done_ = true;
        }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
        return ok_;
    }

private:
    struct Job { std::string name; std::shared_ptr<Image> img; bool ppm; };
    void run() {
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [this] { return done_ || !q_.empty(); });
                if (q_.empty()) return;
                j = std::move(q_.front());
                q_.pop_front();
            }
            bool ok = j.img->savePNG(j.name + ".png");
            if (ok && j.ppm) ok = j.img->savePPM(j.name + ".ppm");
            if (!ok) { std::lock_guard<std::mutex> lk(m_); ok_ = false; }
        }
    }
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<Job> q_;
    std::vector<std::thread> workers_;
    bool done_ = false, ok_ = true;
};
}  // namespace

static char* getCmdOption(char** begin, char** end, const std::string& option) {
    for (char** it = begin; it != end; ++it)
        if (option == *it && it + 1 != end) return *(it + 1);
    return nullptr;
}
static bool cmdOptionExists(char** begin, char** end, const std::string& option) {
    for (char** it = begin; it != end; ++it)
        if (option == *it) return true;
    return false;
}
static void usage() {
    std::cout << "Usage: neuralSDFRenderer [OPTION]... -i SOURCE.h5 [-o DEST_PREFIX]\n"
                 "Options\n"
                 "\t-i input neuralGeometry path (string) REQUIRED\n"
                 "\t-o output image prefix (string) default: {inputPath}; <basename>.png is appended\n"
                 "\t-H imageH (int)  -W imageW (int)   default 512\n"
                 "\t-M matcap file path (string); without it pixels are shaded by the facing ratio\n"
                 "\t-rx -ry -rz rotation in degrees, -z zoom (camera distance, default 2)\n"
                 "\t--single one frame (default here: there is no interactive mode)\n"
                 "\t--spin 360 frames (or --views N) rotating about y\n"
                 "\t--animation 4-input (xyzt) networks\n"
                 "\t--ppm also write <name>.ppm   --scene manysphere|tanh|round|displace|sphere|cylindercut|spherecut   --math reference|deterministic   --max-steps N   --precision N\n"
                 "\t--gpu K   --gpus N (tiles of one frame / views of a turntable over N devices)   --batch B   --stats   -h (--help)\n";
}
static int countDigit(int n) {
    int c = 0;
    do { ++c; n /= 10; } while (n);
    return c;
}

int main(int argc, char** argv) {
    char** end = argv + argc;
    if (cmdOptionExists(argv, end, "-h") || cmdOptionExists(argv, end, "--help")) { usage(); return 0; }
    if (!cmdOptionExists(argv, end, "-i") || !getCmdOption(argv, end, "-i")) {
        std::cerr << "You must give path to neuralGeometry\n";
        return 0;
    }
    const std::string path = getCmdOption(argv, end, "-i");
    const std::string prefix = getCmdOption(argv, end, "-o") ? getCmdOption(argv, end, "-o") : path;
    const int height = getCmdOption(argv, end, "-H") ? atoi(getCmdOption(argv, end, "-H")) : 512;
    const int width = getCmdOption(argv, end, "-W") ? atoi(getCmdOption(argv, end, "-W")) : 512;
    const float rx = getCmdOption(argv, end, "-rx") ? (float)atof(getCmdOption(argv, end, "-rx")) : 0.0f;
    const float ry = getCmdOption(argv, end, "-ry") ? (float)atof(getCmdOption(argv, end, "-ry")) : 0.0f;
    const float zoom = getCmdOption(argv, end, "-z") ? -(float)atof(getCmdOption(argv, end, "-z")) : -2.0f;
    const bool doSpin = cmdOptionExists(argv, end, "--spin") || cmdOptionExists(argv, end, "--views");
    const int views = getCmdOption(argv, end, "--views") ? atoi(getCmdOption(argv, end, "--views")) : 360;
    const int numInputs = cmdOptionExists(argv, end, "--animation") ? 4 : 3;
    const bool ppm = cmdOptionExists(argv, end, "--ppm");
    const int maxSteps = getCmdOption(argv, end, "--max-steps") ? atoi(getCmdOption(argv, end, "--max-steps")) : 6000;
    const char* sceneName = getCmdOption(argv, end, "--scene");
    // default = what the reference's HEAD build renders: sceneSDF returns manySphere(p, nSDF, true) (src/volumeRender_kernel.cu:221)
    const int scene = (sceneName && std::strcmp(sceneName, "tanh") == 0) ? NSDF_SCENE_TANH : NSDF_SCENE_MANYSPHERE;
    const char* mathName = getCmdOption(argv, end, "--math");
    const int mathMode = (mathName && std::strcmp(mathName, "deterministic") == 0) ? NSDF_MATH_DETERMINISTIC : NSDF_MATH_REFERENCE;
    // the scenes sceneSDF (src/volumeRender_kernel.cu:217-230) holds commented out, as scene programs
    nsdf_scene_op prog[2] = {};
    int progOps = 0;
    if (sceneName && std::strcmp(sceneName, "round") == 0) {                 // sdfOpRound(tanh(nSDF), 0.04)
        prog[0].prim = NSDF_PRIM_TANH; prog[1].prim = NSDF_MOD_ROUND; prog[1].a[0] = 0.04f; progOps = 2;
    } else if (sceneName && std::strcmp(sceneName, "displace") == 0) {       // displacementPattern(p, nSDF)
        prog[0].prim = NSDF_PRIM_TANH; prog[1].prim = NSDF_MOD_DISPLACE; prog[1].a[0] = 5.0f; prog[1].a[1] = 0.05f; progOps = 2;
    } else if (sceneName && std::strcmp(sceneName, "sphere") == 0) {         // sdfSphere(p, 0.9)
        prog[0].prim = NSDF_PRIM_SPHERE; prog[0].a[3] = 0.9f; progOps = 1;
    } else if (sceneName && std::strcmp(sceneName, "cylindercut") == 0) {    // manyCylinderCut(p, nSDF)
        prog[0].prim = NSDF_PRIM_NSDF; prog[1].prim = NSDF_MOD_MANY_CYLINDER_CUT; progOps = 2;
    } else if (sceneName && std::strcmp(sceneName, "spherecut") == 0) {      // manySphere(p, nSDF, false)
        prog[0].prim = NSDF_PRIM_NSDF; prog[1].prim = NSDF_MOD_MANY_SPHERE; prog[1].a[0] = 0.0f; progOps = 2;
    } else if (sceneName && std::strcmp(sceneName, "tanh") != 0 && std::strcmp(sceneName, "manysphere") != 0) {
        std::cerr << "unknown --scene " << sceneName << "\n";
        return 1;
    }
    if (getCmdOption(argv, end, "--gpu")) nsdfCheck(nsdf_set_device(atoi(getCmdOption(argv, end, "--gpu"))));
    if (width <= 0 || height <= 0) { std::cerr << "bad image size\n"; return 1; }

    NeuralNetwork nn;
    bool ok = false;
    try {
        ok = nn.load(path);
    } catch (const std::exception& e) {
        std::cerr << e.what() << std::endl;
    }
    if (!ok) { std::cerr << "Failed to load model at " << path << std::endl; return 1; }
    if (getCmdOption(argv, end, "--precision")) nn.setPrecision(atoi(getCmdOption(argv, end, "--precision")));

    Image matcap;
    int colorType = NSDF_COLOR_FACING;          // src/main.cpp:655-662: no -M means facing-ratio shading
    if (getCmdOption(argv, end, "-M")) {
        if (!matcap.loadPNG(getCmdOption(argv, end, "-M"))) return 1;
        colorType = NSDF_COLOR_MATCAP;
    }
    copyStaticSettings(colorType, numInputs);
    setRenderOptions(scene, maxSteps, 0);
    setMathMode(mathMode);
    if (progOps) setSceneProgram(prog, progOps);

    const int gpus = getCmdOption(argv, end, "--gpus") ? atoi(getCmdOption(argv, end, "--gpus")) : 1;
    int batch = getCmdOption(argv, end, "--batch") ? atoi(getCmdOption(argv, end, "--batch")) : 16;
    if (batch < 1) batch = 1;
    if ((long long)width * height > (1LL << 24)) batch = 1;     // nsdf_render_views: batches need W * H <= 2^24
    if (gpus < 1 || gpus > nsdf_device_count()) { std::cerr << "--gpus: " << nsdf_device_count() << " device(s) present\n"; return 1; }
    const int frames = doSpin ? views : 1;
    const int per_round = doSpin ? std::min(frames, batch * gpus) : 1;      // frames rendered per call
    if (per_round > 128 * gpus) { std::cerr << "--batch: at most 128 views per device and launch\n"; return 1; }

    nsdf_model* model = nn.handle();
    if (!model) { std::cerr << "the network has no dense layers\n"; return 1; }
    nsdf_multi* multi = nullptr;
    if (gpus > 1) {
        std::vector<int> devs;
        int first = 0;
        if (getCmdOption(argv, end, "--gpu")) first = atoi(getCmdOption(argv, end, "--gpu"));
        for (int i = 0; i < gpus; ++i) devs.push_back((first + i) % nsdf_device_count());
        nsdfCheck(nsdf_multi_create(model, devs.data(), gpus, &multi));
    }
    uint* d_output = nullptr;
    const size_t frame_px = (size_t)width * height;
    nsdfCheck(nsdf_dev_malloc(reinterpret_cast<void**>(&d_output), frame_px * per_round * sizeof(uint)));
    nsdf_render_stats* d_stats = nullptr;
    nsdfCheck(nsdf_dev_malloc(reinterpret_cast<void**>(&d_stats), sizeof(nsdf_render_stats)));
    const dim3 blockSize(8, 8);
    const std::string base = path.substr(path.find_last_of("/\\") + 1);
    nsdf_render_opts opts;
    getRenderOptions(&opts);
    const bool useMatcap = colorType == NSDF_COLOR_MATCAP;
    unsigned hw = std::thread::hardware_concurrency();
    FrameWriter writer((int)std::max(1u, std::min(8u, hw ? hw : 1u)));

    for (int i0 = 0; i0 < frames; i0 += per_round) {
        const int n = std::min(per_round, frames - i0);
        std::vector<nsdf_camera> cams((size_t)n);
        for (int k = 0; k < n; ++k) {
            const int i = i0 + k;
            const float yaw = doSpin ? ry + 360.0f * i / views : ry;     // doABarrelRoll: viewRotation.y = i, frameNumber = i
            nsdf_camera_from_view(rx, yaw, 0.0f, 0.0f, zoom, doSpin ? i : 0, &cams[(size_t)k]);
        }
        nsdfCheck(nsdf_dev_memset(d_output, 0, frame_px * n * sizeof(uint)));     // generateSingleImage: cudaMemset (main.cpp:407-408)
        nsdf_render_stats st;
        std::memset(&st, 0, sizeof(st));
        nsdfCheck(nsdf_device_synchronize());
        const auto t0 = std::chrono::steady_clock::now();
        if (multi) {
            nsdfCheck(nsdf_multi_render(multi, cams.data(), n, useMatcap ? matcap.hostData.get() : nullptr, (int)matcap.shape.x,
                                        (int)matcap.shape.y, &opts, d_output, width, height, &st));
        } else {
            nsdfCheck(nsdf_dev_memset(d_stats, 0, sizeof(nsdf_render_stats)));
            nsdfCheck(nsdf_render_views(model, cams.data(), n, useMatcap ? matcap.deviceData.get() : nullptr, (int)matcap.shape.x,
                                        (int)matcap.shape.y, &opts, d_output, width, height, d_stats, nullptr));
            nsdfCheck(nsdf_device_synchronize());
            nsdfCheck(nsdf_memcpy_d2h(&st, d_stats, sizeof(st)));
        }
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (i0 == 0)
            std::printf("volumeRender, Throughput = %.4f MTexels/s, Time = %.5f s, Size = %u Texels, NumDevsUsed = %u, Workgroup = %u\n",
                        (1.0e-6 * width * height * n) / dt, dt / n, (unsigned)(width * height), (unsigned)gpus, blockSize.x * blockSize.y);
        if (cmdOptionExists(argv, end, "--stats"))
            std::printf("frame %d%s: %.3f ms%s, %llu MLP queries (%.2f Gq/s), %llu hit pixels, %llu rays, %llu unfinished\n", i0,
                        n > 1 ? (".." + std::to_string(i0 + n - 1)).c_str() : "", dt * 1e3 / n, n > 1 ? " per frame" : "", st.queries,
                        st.queries / dt * 1e-9, st.hit_pixels, st.rays, st.unfinished);
        for (int k = 0; k < n; ++k) {
            const int i = i0 + k;
            auto out = std::make_shared<Image>(width, height, true);
            out->allocateMemory();
            nsdfCheck(nsdf_memcpy_d2h(out->hostData.get(), d_output + (size_t)k * frame_px, frame_px * sizeof(uint)));
            std::string name;
            if (doSpin) {
                if (countDigit(i) < 2) name = "00";
                else if (countDigit(i) == 2) name = "0";
                name += std::to_string(i);
            } else {
                name = base;
            }
            std::cout << "saving frame: " << prefix + name + ".png" << std::endl;
            writer.push(prefix + name, out, ppm);      // encoded and written by a host thread while the next batch renders
        }
    }
    const bool wrote = writer.finish();
    if (multi) nsdf_multi_destroy(multi);
    nsdf_dev_free(d_stats);
    nsdf_dev_free(d_output);
    return wrote ? 0 : 1;
}
