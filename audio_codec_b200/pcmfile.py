"""Drop-in for `pcmfile.PCMFile` (coder/pcmfile.py:29-147): 16-bit PCM WAV reader / writer.
File I/O is host glue; the int16 <-> signed-fraction mapping goes through the same
uniform quantiser kernels as the reference goes through vQuantizeUniform /
vDequantizeUniform (so -32768 reads as 0.0 and full scale is +-32767/65535*2)."""
from struct import calcsize, pack, unpack

import numpy as np

from .audiofile import AudioFile, CodingParams
from .quantize import vDequantizeUniform, vQuantizeUniform

__all__ = ["PCMFile"]


class PCMFile(AudioFile):

    def ReadFileHeader(self):
        """RIFF/WAVE header -> CodingParams(nChannels, bitsPerSample, sampleRate, numSamples)
        (coder/pcmfile.py:32-64)."""
        tag = self.fp.read(12)
        if tag[0:4] != b"RIFF" or tag[8:12] != b"WAVE":
            raise RuntimeError("ERROR: File opened for PCMFile is not a RIFF file!")
        while True:
            tag = self.fp.read(4)
            if len(tag) < 4:
                raise RuntimeError("ERROR: Didn't find WAV file 'fmt ' chunk following RIFF file header")
            if tag == b"fmt ":
                break
        formatSize, formatTag, nChannels, sampleRate, _, _, bitsPerSample = unpack("<LHHLLHH", self.fp.read(20))
        if formatTag != 1:
            raise IOError("Opened a non-PCM WAV file as a PCMFile")
        if bitsPerSample != 16:
            raise RuntimeError("PCMFile was not 16-bits per sample")
        while True:
            tag = self.fp.read(4)
            if len(tag) < 4:
                raise RuntimeError("Didn't find WAV file 'data' chunk following 'fmt ' chunk")
            if tag == b"data":
                break
        numBytes = unpack("<L", self.fp.read(4))[0]
        p = CodingParams()
        p.nChannels = nChannels
        p.bitsPerSample = bitsPerSample
        p.sampleRate = sampleRate
        p.numSamples = numBytes // (nChannels * (bitsPerSample // 8))
        p.bytesReadSoFar = 0
        return p

    def ReadDataBlock(self, codingParams):
        """Next nSamplesPerBlock samples per channel as signed fractions; the final partial block is
        zero padded; None at the end (coder/pcmfile.py:66-100)."""
        cp = codingParams
        bps = cp.bitsPerSample // 8
        want = cp.nSamplesPerBlock * cp.nChannels * bps
        left = cp.nChannels * cp.numSamples * bps - cp.bytesReadSoFar
        raw = None if left <= 0 else self.fp.read(min(left, want))
        cp.bytesReadSoFar += want
        if not raw:
            return None
        raw = raw + b"\0" * (want - len(raw))
        codes = np.frombuffer(raw, dtype="<i2").astype(np.int64).reshape(-1, cp.nChannels).T
        neg = codes < 0
        mag = vDequantizeUniform(np.ascontiguousarray(np.abs(codes)), 16)      # -32768 -> magnitude bits 0 -> 0.0
        return [np.where(neg[c], -mag[c], mag[c]) for c in range(cp.nChannels)]

    def WriteFileHeader(self, codingParams):
        """(coder/pcmfile.py:103-115)"""
        cp = codingParams
        bps = cp.bitsPerSample // 8
        dataBytes = cp.numSamples * cp.nChannels * bps
        self.fp.write(pack("<4sL4s4sLHHLLHH4sL", b"RIFF", 36 + dataBytes, b"WAVE", b"fmt ", 16, 1, cp.nChannels,
                           cp.sampleRate, cp.sampleRate * cp.nChannels * bps, cp.nChannels * bps, cp.bitsPerSample,
                           b"data", dataBytes))

    def WriteDataBlock(self, data, codingParams):
        """Signed fractions -> interleaved 16-bit codes (coder/pcmfile.py:118-147)."""
        if len(data) != codingParams.nChannels:
            raise RuntimeError("Data block to PCMFile did not have expected number of channels")
        n = min(len(d) for d in data)
        x = np.stack([np.asarray(d, dtype=np.float64)[:n] for d in data])
        neg = np.signbit(x)
        codes = vQuantizeUniform(np.abs(x), 16).astype(np.int16)
        codes = np.where(neg, -codes, codes).astype("<i2")
        self.fp.write(np.ascontiguousarray(codes.T).tobytes())
