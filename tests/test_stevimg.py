"""`.stevimg` files (io/image_io.h:29-165): the Python module and the C++ facade header against the
reference's own writer and reader (compiled into oracle/_ref/libstevi_ref_ibr.so) — files written by one side are read by the
others with the same shape, strides and values, including the reference's cost-volume layout {W*D, 1, W}, type conversion on
write, missing trailing dimensions and the refusals (wrong type, too many dimensions, missing file)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import oracle as O
from libstevi_b200 import stevimg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_ref = pytest.mark.skipif(O.ref_ibr is None or not O.ref_ibr.has("write_stevimg_f32"), reason="oracle/_ref/libstevi_ref_ibr.so was not built")


def test_descriptors():
    assert [stevimg.dtypeDescr(t) for t in (np.uint8, np.uint16, np.int32, np.float32, np.float64)] == ["u8", "u16", "i32", "f32", "f64"]


def test_python_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    vol = rng.uniform(-1, 1, (5, 7, 3)).astype(np.float32)
    ref_layout = np.lib.stride_tricks.as_strided(np.ascontiguousarray(vol.transpose(0, 2, 1)), shape=vol.shape, strides=(7 * 3 * 4, 4, 7 * 4))
    assert np.array_equal(ref_layout, vol)
    for name, arr in (("c.stevimg", vol), ("ref.stevimg", ref_layout), ("sliced.stevimg", vol[:, ::2])):
        path = str(tmp_path / name)
        assert stevimg.writeStevimg(path, arr)
        back = stevimg.readStevimg(path, np.float32, 3)
        assert back.shape == arr.shape and np.array_equal(back, arr)
    assert tuple(s // 4 for s in stevimg.readStevimg(str(tmp_path / "ref.stevimg"), np.float32, 3).strides) == (21, 1, 7)   # kept as written
    path = str(tmp_path / "disp.stevimg")
    disp = rng.integers(0, 200, (6, 9)).astype(np.int32)
    assert stevimg.writeStevimg(path, disp, np.uint8)   # conversion on write (:70-82)
    assert np.array_equal(stevimg.readStevimg(path, np.uint8, 2), disp.astype(np.uint8))
    assert stevimg.readStevimg(path, np.uint8, 3).shape == (6, 9, 1)   # missing dimensions are 1 (:143-156)
    assert stevimg.readStevimg(path, np.float32, 2) is None and stevimg.readStevimg(path, np.uint8, 1) is None
    assert stevimg.readStevimg(str(tmp_path / "absent.stevimg"), np.uint8, 2) is None
    assert stevimg.stevImgFileMatchTypeAndDim(path, np.uint8, 2) and not stevimg.stevImgFileMatchTypeAndDim(path, np.float32, 2)


@needs_ref
def test_files_of_the_reference_and_ours_are_interchangeable(tmp_path):
    lib = O.ref_ibr.lib
    rng = np.random.default_rng(1)
    vol = rng.uniform(-1, 1, (4, 6, 5)).astype(np.float32)
    for ref_layout in (0, 1):
        path = str(tmp_path / f"ref{ref_layout}.stevimg").encode()
        assert lib.ref_write_stevimg_f32(path, vol.ctypes.data_as(ctypes.c_void_p), 4, 6, 5, ref_layout) == 0
        back = stevimg.readStevimg(path.decode(), np.float32, 3)
        assert np.array_equal(back, vol)
        if ref_layout:
            assert tuple(s // 4 for s in back.strides) == (30, 1, 6)
        # and ours, read by the reference
        mine = str(tmp_path / f"mine{ref_layout}.stevimg")
        assert stevimg.writeStevimg(mine, back)
        assert open(mine, "rb").read() == open(path.decode(), "rb").read()   # byte for byte the reference's file
        out = np.zeros_like(vol)
        shape = (ctypes.c_int * 3)()
        assert lib.ref_read_stevimg_f32(mine.encode(), out.ctypes.data_as(ctypes.c_void_p), shape) == 0
        assert tuple(shape) == vol.shape and np.array_equal(out, vol)
    disp = rng.integers(0, 250, (5, 8)).astype(np.int32)
    path = str(tmp_path / "disp.stevimg")
    assert lib.ref_write_stevimg_u8_from_i32(path.encode(), disp.ctypes.data_as(ctypes.c_void_p), 5, 8) == 0
    mine = str(tmp_path / "disp_mine.stevimg")
    assert stevimg.writeStevimg(mine, disp, np.uint8)
    # the converted copy takes the array library's default strides (a fiat of the stand-in): same values, any order
    assert np.array_equal(stevimg.readStevimg(path, np.uint8, 2), stevimg.readStevimg(mine, np.uint8, 2))
    assert np.array_equal(stevimg.readStevimg(path, np.uint8, 2), disp.astype(np.uint8))
    assert lib.ref_stevimg_matches(mine.encode(), 1) == 1 and lib.ref_stevimg_matches(mine.encode(), 2) == 0


def test_facade_header_reads_and_writes(tmp_path):
    """include/StereoVision/io/image_io.h, compiled with plain g++: writes a strided volume and a converted map, reads ours back."""
    rng = np.random.default_rng(2)
    vol = rng.uniform(-1, 1, (3, 5, 4)).astype(np.float32)
    src_path = str(tmp_path / "from_python.stevimg")
    assert stevimg.writeStevimg(src_path, vol)
    cpp = os.path.join(str(tmp_path), "io.cpp")
    open(cpp, "w").write(r'''
#include <cstdio>
#include <StereoVision/io/image_io.h>
using namespace StereoVision;
int main(int argc, char** argv) {
	auto a = IO::readStevimg<float, 3>(argv[1]);
	if (a.empty()) return 2;
	Multidim::Array<float, 3> b({a.shape()[0], a.shape()[1], a.shape()[2]}, {a.shape()[1] * a.shape()[2], 1, a.shape()[1]});   // the reference's volume layout
	for (int i = 0; i < a.shape()[0]; i++) for (int j = 0; j < a.shape()[1]; j++) for (int k = 0; k < a.shape()[2]; k++) b.atUnchecked(i, j, k) = 2 * a.valueUnchecked(i, j, k);
	if (!IO::writeStevimg<float, float, 3>(argv[2], b)) return 3;
	if (!IO::writeStevimg<uint8_t, float, 3>(argv[3], a)) return 4;
	std::printf("%d %d %d %d\n", int(IO::stevImgFileMatchTypeAndDim<float, 3>(argv[1])), int(IO::stevImgFileMatchTypeAndDim<uint8_t, 3>(argv[1])),
				int(IO::stevImgFileMatchTypeAndDim<float, 2>(argv[1])), int(IO::readStevimg<float, 2>(argv[1]).empty()));
	return 0;
}''')
    exe = os.path.join(str(tmp_path), "io")
    r = subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), cpp, "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    doubled, as_u8 = str(tmp_path / "doubled.stevimg"), str(tmp_path / "u8.stevimg")
    r = subprocess.run([exe, src_path, doubled, as_u8], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.split() == ["1", "0", "0", "1"]
    back = stevimg.readStevimg(doubled, np.float32, 3)
    assert np.array_equal(back, 2 * vol) and tuple(s // 4 for s in back.strides) == (20, 1, 5)
    assert np.array_equal(stevimg.readStevimg(as_u8, np.uint8, 3), vol.astype(np.uint8))
