"""CPU tests of the C++ host layer (include/*.hh, hostcpp/, apps/): the reference-compatible classes over the
C ABI.  No GPU: only host-side behaviour (hostOnly loads, containers, PNG round trip) and that the
reference's own translation units compile against these headers unmodified."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import MATCAPS, ROOT, model_path

INC = os.path.join(ROOT, "include")
LIBDIR = os.path.join(ROOT, "cudaneuralrender_b200")
REF = "/root/reference/src"


def _compile(tmp_path, name, code, extra=()):
    src = tmp_path / (name + ".cpp")
    src.write_text(code)
    exe = tmp_path / name
    cmd = ["g++", "-std=c++17", "-Wall", "-Werror", "-I" + INC, str(src), *extra, "-o", str(exe), "-L" + LIBDIR, "-lnsdf",
           "-Wl,-rpath," + LIBDIR]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return str(exe)


def test_apps_are_built(built):
    for exe in ("neuralSDFRenderer", "simpleInfer"):
        assert os.access(os.path.join(ROOT, "apps", exe), os.X_OK)
    out = subprocess.run([os.path.join(ROOT, "apps", "neuralSDFRenderer"), "-h"], capture_output=True, text=True)
    assert out.returncode == 0 and "-i input neuralGeometry path" in out.stdout
    out = subprocess.run([os.path.join(ROOT, "apps", "neuralSDFRenderer")], capture_output=True, text=True)
    assert "You must give path to neuralGeometry" in out.stderr


def test_host_classes_load_and_layout(built, tmp_path):
    """hostOnly load (no device): layer order, [out][in] weight layout, parameter counts, Matrix semantics."""
    exe = _compile(tmp_path, "t_load", r'''
#include "neuralNetwork.hh"
#include "layers/denseLayer.hh"
#include <cstdio>
int main(int, char** argv) {
    NeuralNetwork nn;
    if (!nn.load(argv[1], true)) return 2;
    std::printf("%zu %d %d\n", nn.getLayers().size(), nn.getNumWeightParams(), nn.getNumBiasParams());
    for (Layer* l : nn.getLayers()) {
        DenseLayer* d = static_cast<DenseLayer*>(l);
        Matrix W = d->getWeightsMatrix(); Matrix b = d->getBiasVector();
        std::printf("%s %d %d %d %.9g %.9g %.9g\n", d->getName().c_str(), d->getXDim(), d->getYDim(), d->getActivation(),
                    W[0], W[W.size() - 1], b[0]);
    }
    Matrix A(Shape(3, 2), true); A.allocateMemory(); A[4] = 7.f;
    Matrix B = A;                       // copies alias the storage (shared_ptr), as in the reference
    B[4] = 9.f;
    std::printf("alias %g %d\n", A[4], A.size());
    return 0;
}
''')
    out = subprocess.run([exe, model_path("car_1")], capture_output=True, text=True, check=True).stdout.split("\n")
    assert out[0].split() == ["9", "7296", "257"]
    from oracle import h5mini
    layers = h5mini.load_keras_mlp(model_path("car_1"))
    for i, (kernel, bias) in enumerate(layers):
        f = out[1 + i].split()
        k = np.asarray(kernel, dtype=np.float32)                # [in][out]
        assert f[0] == f"Dense_{i}" and int(f[1]) == k.shape[0] and int(f[2]) == k.shape[1]
        assert int(f[3]) == (1 if i == len(layers) - 1 else 0)  # last layer "Tanh", the others ReLU
        assert np.float32(f[4]) == k[0, 0] and np.float32(f[5]) == k[-1, -1]   # W[y*in + x] = kernel[x][y]
        assert np.float32(f[6]) == np.float32(np.asarray(bias).ravel()[0])
    assert out[10].split() == ["alias", "9", "6"]


def test_load_errors_like_reference(built, tmp_path):
    """Missing file -> exception (HighFive behaviour); path constructor -> exit(1) (neuralNetwork.cpp:36-42)."""
    exe = _compile(tmp_path, "t_err", r'''
#include "neuralNetwork.hh"
#include <cstdio>
int main(int argc, char** argv) {
    if (argc > 2) { NeuralNetwork nn(argv[1]); return 0; }
    try { NeuralNetwork nn; nn.load(argv[1], true); } catch (const HighFive::Exception& e) { std::printf("caught\n"); return 0; }
    return 3;
}
''')
    bad = tmp_path / "garbage.h5"
    bad.write_bytes(b"not an hdf5 file at all" * 10)
    assert "caught" in subprocess.run([exe, "/nonexistent/model.h5"], capture_output=True, text=True).stdout
    assert "caught" in subprocess.run([exe, str(bad)], capture_output=True, text=True).stdout


def test_image_png_round_trip(built, tmp_path):
    exe = _compile(tmp_path, "t_img", r'''
#include "neuralUtils/image.hh"
#include <cstdio>
int main(int, char** argv) {
    int w = 0, h = 0;
    if (nsdf_png_load(argv[1], nullptr, 0, &w, &h) != NSDF_OK) return 2;
    Image im(w, h, true); im.allocateMemory();
    if (nsdf_png_load(argv[1], im.hostData.get(), (size_t)w * h, &w, &h) != NSDF_OK) return 3;
    if (!im.savePNG(argv[2], false, false) || !im.savePNG(argv[3]) || !im.savePPM(argv[4])) return 4;
    std::printf("%d %d %u\n", w, h, im[w + 1]);
    return 0;
}
''')
    src = os.path.join(MATCAPS, "ball-298x300.png")
    a, b, c = (str(tmp_path / n) for n in ("same.png", "turned.png", "out.ppm"))
    out = subprocess.run([exe, src, a, b, c], capture_output=True, text=True, check=True).stdout.split()
    from PIL import Image as PImage
    ref = np.asarray(PImage.open(src).convert("RGBA"))
    assert (int(out[0]), int(out[1])) == (ref.shape[1], ref.shape[0])
    assert np.array_equal(np.asarray(PImage.open(a).convert("RGBA")), ref)
    # the reference's default savePNG pushes a,b,g,r per texel and reverses the byte stream (image.cu:84-98):
    # the texel ORDER is reversed (a 180 degree turn), the channels come out as r,g,b,a again
    turned = np.asarray(PImage.open(b).convert("RGBA"))
    assert np.array_equal(turned, ref[::-1, ::-1, :])
    ppm = open(c, "rb").read()
    head = f"P6\n{ref.shape[1]}\n{ref.shape[0]}\n255\n".encode()
    assert ppm.startswith(head) and ppm[len(head):] == ref[:, :, :3].tobytes()


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sources are only present in the build container")
def test_reference_sources_compile_against_these_headers(built, tmp_path):
    """API compatibility: the reference's own simpleInfer.cpp (with its private copy of the loader) and
    neuralNetwork.cpp compile UNMODIFIED against include/ (copied next to nothing of the reference, so their
    quoted includes resolve to this repository's headers) and link against libnsdf.so."""
    for f in ("simpleInfer.cpp", "neuralNetwork.cpp"):
        shutil.copy(os.path.join(REF, f), tmp_path / f)
        os.chmod(tmp_path / f, 0o644)
    r = subprocess.run(["g++", "-std=c++14", "-I" + INC, str(tmp_path / "simpleInfer.cpp"), "-o", str(tmp_path / "simpleInfer"),
                        "-L" + LIBDIR, "-lnsdf", "-Wl,-rpath," + LIBDIR], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run(["g++", "-std=c++14", "-c", "-DNSDF_REFERENCE_NEURALNETWORK_CPP", "-I" + INC, str(tmp_path / "neuralNetwork.cpp"),
                        "-o", str(tmp_path / "nn.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    syms = subprocess.run(["nm", "-C", str(tmp_path / "nn.o")], capture_output=True, text=True).stdout
    assert "NeuralNetwork::load(" in syms and "NeuralNetwork::forward(Matrix, int)" in syms
