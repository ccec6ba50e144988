"""``.stevimg`` files (LibStevi's io/image_io.h:29-165): one text line ``<dtype> <nDim> <shape...> <strides...>`` — dtype = i|u|f +
bit count (utils/types_manipulations.h:82-102), strides in elements — then the elements as they lie in memory. Host-side file io
for exchanging images, disparity maps and volumes with LibStevi installs; results come back as numpy views with the file's own
strides, so a cost volume written by the reference in its {W*D, 1, W} layout is read without a copy."""
import numpy as np

_DESCR = {"u8": np.uint8, "u16": np.uint16, "u32": np.uint32, "u64": np.uint64, "i8": np.int8, "i16": np.int16, "i32": np.int32,
          "i64": np.int64, "f32": np.float32, "f64": np.float64}


def dtypeDescr(dtype):
    """TypesManipulations::dtypeDescr<T>(), utils/types_manipulations.h:82-102."""
    dt = np.dtype(dtype)
    kind = {"u": "u", "i": "i", "f": "f", "b": "u"}.get(dt.kind)
    return "" if kind is None else f"{kind}{dt.itemsize * 8}"


def _dense(strides, shape):
    """Does (shape, strides) address a gap-free block of prod(shape) elements starting at element 0 (Multidim::Array::isDense)?"""
    n = int(np.prod(shape)) if len(shape) else 1
    if n == 0:
        return True
    if any(s < 0 for s in strides):
        return False
    return 1 + sum((d - 1) * s for d, s in zip(shape, strides)) == n


def writeStevimg(fileName, image, ImgType=None):
    """writeStevimg<ImgType>(fileName, image), io/image_io.h:66-127: converts to ImgType, makes the array dense when it is not,
    writes the header and the elements in memory order. Returns False when the file cannot be written."""
    a = np.asarray(image)
    if ImgType is not None and a.dtype != np.dtype(ImgType):
        a = a.astype(ImgType)
    if dtypeDescr(a.dtype) not in _DESCR:
        return False
    strides = [s // a.itemsize for s in a.strides]
    if not _dense(strides, a.shape) or any(s % a.itemsize for s in a.strides):
        a = np.ascontiguousarray(a)
        strides = [s // a.itemsize for s in a.strides]
    head = " ".join([dtypeDescr(a.dtype), str(a.ndim)] + [str(int(v)) for v in a.shape] + [str(int(v)) for v in strides]) + "\n"
    try:
        with open(fileName, "wb") as f:
            f.write(head.encode("ascii"))
            if a.size:
                # the elements as they lie in memory: view the block through its own strides as a flat run
                flat = np.lib.stride_tricks.as_strided(a, shape=(a.size,), strides=(a.itemsize,))
                f.write(flat.tobytes())
        return True
    except OSError:
        return False


def stevImgFileMatchTypeAndDim(fileName, ImgType, nDim):
    """io/image_io.h:29-64."""
    try:
        with open(fileName, "rb") as f:
            parts = f.readline().split()
        return len(parts) >= 2 and parts[0].decode() == dtypeDescr(ImgType) and int(parts[1]) <= nDim
    except (OSError, ValueError):
        return False


def readStevimg(fileName, ImgType, nDim):
    """readStevimg<ImgType, nDim>(fileName), io/image_io.h:129-165. None where the reference returns an empty array (missing file,
    another element type, more dimensions than asked for); missing trailing dimensions have extent 1."""
    try:
        with open(fileName, "rb") as f:
            parts = f.readline().split()
            if len(parts) < 2 or parts[0].decode() != dtypeDescr(ImgType):
                return None
            n = int(parts[1])
            if n > nDim or len(parts) < 2 + 2 * n:
                return None
            shape = [int(v) for v in parts[2:2 + n]] + [1] * (nDim - n)
            strides = [int(v) for v in parts[2 + n:2 + 2 * n]] + [1] * (nDim - n)
            count = int(np.prod(shape))
            data = np.frombuffer(f.read(count * np.dtype(ImgType).itemsize), dtype=ImgType).copy()
    except (OSError, ValueError):
        return None
    if data.size != count:
        return None
    item = data.itemsize
    return np.lib.stride_tricks.as_strided(data, shape=shape, strides=[s * item for s in strides])
