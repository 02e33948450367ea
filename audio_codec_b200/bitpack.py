"""Drop-in for `bitpack.PackedBits` (coder/bitpack.py:13-175): an MSB-first bit
container.  Packing and unpacking run on the GPU (one warp per row: exclusive scan
of the field widths, then parallel writers); WriteBits queues fields and the queue
is packed in one launch when the bytes are needed.  Fields may be wider than 64
bits (Python ints), as in the reference."""
import numpy as np
import torch

from . import _lib
from ._util import ptr

__all__ = ["PackedBits", "BYTESIZE", "pack_rows", "unpack_rows"]

BYTESIZE = 8


def _split(value, width):
    """(value, width) -> chunks of at most 64 bits, most significant first."""
    value = int(value) & ((1 << width) - 1) if width > 0 else 0
    out = []
    while width > 64:
        width -= 64
        out.append(((value >> width) & 0xFFFFFFFFFFFFFFFF, 64))
    out.append((value & ((1 << width) - 1) if width else 0, width))
    return out


def pack_rows(values, widths, counts, row_bytes):
    """Batched packer: values/widths [rows, max_fields] (uint64 as int64 bit patterns / int32),
    counts [rows] -> uint8 CUDA tensor [rows, row_bytes]."""
    dev = torch.device("cuda", torch.cuda.current_device())
    v = torch.as_tensor(values, device=dev).contiguous()
    w = torch.as_tensor(widths, dtype=torch.int32, device=dev).contiguous()
    c = torch.as_tensor(counts, dtype=torch.int32, device=dev).contiguous()
    rows, maxf = w.shape
    row_bytes = (int(row_bytes) + 3) // 4 * 4
    out = torch.empty((rows, row_bytes), dtype=torch.uint8, device=dev)
    err = _lib.C.c_int(0)
    rc = _lib.lib.pac_pack_bits(_lib.default_context().h, ptr(v), ptr(w), ptr(c), maxf, rows, row_bytes, ptr(out),
                                _lib.C.byref(err), _lib.stream_ptr())
    _lib.check_soft(rc, err)
    return out


def unpack_rows(data, start_bits, widths, counts):
    """Batched reader: data uint8 [rows, row_bytes]; fields of row r start at start_bits[r]."""
    dev = torch.device("cuda", torch.cuda.current_device())
    d = torch.as_tensor(data, dtype=torch.uint8, device=dev).contiguous()
    w = torch.as_tensor(widths, dtype=torch.int32, device=dev).contiguous()
    c = torch.as_tensor(counts, dtype=torch.int32, device=dev).contiguous()
    s = torch.as_tensor(start_bits, dtype=torch.int64, device=dev).contiguous()
    rows, maxf = w.shape
    out = torch.zeros((rows, maxf), dtype=torch.int64, device=dev)
    _lib.check(_lib.lib.pac_unpack_bits(_lib.default_context().h, ptr(d), d.shape[1], ptr(s), ptr(w), ptr(c), maxf, rows,
                                        ptr(out), _lib.stream_ptr()))
    return out


class PackedBits:
    """Array of bytes read and written as individual bits, MSB first."""

    def __init__(self):
        self.iByte = self.iBit = 0
        self.nBytes = 0
        self._data = np.zeros(0, dtype=np.uint8)
        self._pending = []          # (value, width) chunks not yet packed
        self._pend_pos = 0          # bit position of the first pending chunk
        self._dev = None            # cached device copy of the bytes

    # ---- cursor helpers
    def _pos(self):
        return self.iByte * 8 + self.iBit

    def _seek(self, pos):
        self.iByte, self.iBit = divmod(int(pos), 8)

    def Size(self, nBytes):
        """Holds nBytes of zeros; resets the cursor (coder/bitpack.py:20-24)."""
        self.nBytes = int(nBytes)
        self.iByte = self.iBit = 0
        self._data = np.zeros(self.nBytes, dtype=np.uint8)
        self._pending = []
        self._dev = None

    def _flush(self):
        if not self._pending:
            return
        start = self._pend_pos
        b0, lead = divmod(start, 8)
        fields = list(self._pending)
        self._pending = []
        if lead:                                   # keep the bits already in the first byte
            fields.insert(0, (int(self._data[b0]) >> (8 - lead), lead))
        total = sum(w for _, w in fields)
        nbytes = (total + 7) // 8
        if b0 + nbytes > self.nBytes:
            raise IndexError("index %d is out of bounds for axis 0 with size %d" % (self.nBytes, self.nBytes))
        vals = np.array([[v - (1 << 64) if v >= (1 << 63) else v for v, _ in fields]], dtype=np.int64)
        wid = np.array([[w for _, w in fields]], dtype=np.int32)
        packed = pack_rows(vals, wid, [len(fields)], nbytes)[0, :nbytes].cpu().numpy()
        seg = self._data[b0:b0 + nbytes]
        if lead:
            seg[0] = 0
        np.bitwise_or(seg, packed, out=seg)        # the reference adds into zero-initialised bytes
        self._dev = None

    @property
    def data(self):
        self._flush()
        return self._data

    @data.setter
    def data(self, v):
        self._pending = []
        self._data = np.array(v, dtype=np.uint8)
        self._dev = None

    def device_data(self):
        """The bytes as a CUDA tensor (cached until the next write)."""
        self._flush()
        if self._dev is None:
            self._dev = torch.from_numpy(np.ascontiguousarray(self._data)).cuda()
        return self._dev

    def GetPackedData(self):
        return self.data.tobytes()

    def SetPackedData(self, data):
        self.nBytes = len(data)
        self._pending = []
        self._data = np.frombuffer(bytes(data), dtype=np.uint8).copy()
        self._dev = None

    def WriteBits(self, info, nBits):
        """Writes the lowest nBits of info at the cursor (coder/bitpack.py:37-102)."""
        nBits = int(nBits)
        if nBits <= 0:
            return
        if not self._pending:
            self._pend_pos = self._pos()
        elif self._pend_pos + sum(w for _, w in self._pending) != self._pos():
            self._flush()
            self._pend_pos = self._pos()
        self._pending.extend(_split(int(info), nBits))
        self._seek(self._pos() + nBits)

    def ReadFields(self, widths):
        """Reads consecutive fields of the given widths (<= 64 bits each) in one launch."""
        widths = [int(w) for w in widths]
        if not widths:
            return []
        d = self.device_data().reshape(1, -1)
        vals = unpack_rows(d, [self._pos()], [widths], [len(widths)])[0].cpu().numpy()
        self._seek(self._pos() + sum(widths))
        return [int(v) & 0xFFFFFFFFFFFFFFFF for v in vals]

    def ReadBits(self, nBits):
        """Returns the next nBits at the cursor as an int (coder/bitpack.py:105-175)."""
        nBits = int(nBits)
        if nBits <= 0:
            return 0
        chunks = [w for _, w in _split(0, nBits)]
        out = 0
        for v, w in zip(self.ReadFields(chunks), chunks):
            out = (out << w) | v
        return out

    def ResetPointers(self):
        self.iByte = self.iBit = 0
