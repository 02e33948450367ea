"""Oracle: MSB-first bit writer / reader.  TEST INFRASTRUCTURE ONLY.

Same byte layout as coder/bitpack.py:37-175 (PackedBits.WriteBits/ReadBits):
fields are written most-significant bit first into a zero-initialised byte
array; only the low `n` bits of a value are kept.
"""
import numpy as np


class BitWriter:
    def __init__(self, nbytes):
        self.buf = bytearray(nbytes)
        self.pos = 0  # bit cursor

    def WriteBits(self, value, n):
        value = int(value) & ((1 << n) - 1) if n > 0 else 0
        for i in range(n - 1, -1, -1):
            if (value >> i) & 1:
                self.buf[self.pos >> 3] |= 0x80 >> (self.pos & 7)
            self.pos += 1

    def bytes(self):
        return bytes(self.buf)


class BitReader:
    def __init__(self, data):
        self.buf = bytes(data)
        self.pos = 0

    def ReadBits(self, n):
        v = 0
        for _ in range(n):
            v = (v << 1) | ((self.buf[self.pos >> 3] >> (7 - (self.pos & 7))) & 1)
            self.pos += 1
        return v
