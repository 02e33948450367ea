"""Drop-in for `pacfile.PACFile` (coder/pacfile.py:128-668) over the device codec.

The block-at-a-time API of the reference is kept: WriteDataBlock encodes the
channels of one block with one launch of the fused frame kernel and writes the
length-prefixed records; ReadDataBlock decodes one block's records with one launch
and overlap-adds.  For throughput use the whole-stream entry points
(`encode_file`, `decode_file`, `audio_codec_b200.engine`), which keep every stage of
every block of every stream on the GPU in a handful of launches.
"""
from struct import calcsize, pack, unpack

import numpy as np
import torch

from . import _lib, codec, engine
from ._util import ptr
from .audiofile import AudioFile, CodingParams
from .psychoac import AssignMDCTLinesFromFreqLimits, ScaleFactorBands
from .sbr import omitted_bands

__all__ = ["PACFile", "encode_file", "decode_file", "encode_decode_test"]


class PACFile(AudioFile):
    tag = b"PAC "

    # ------------------------------------------------------------------ headers
    def ReadFileHeader(self):
        """(coder/pacfile.py:136-175); also the 7-field header of the test_decoded/ generation (no useVQ field:
        block floating point), see engine.parse_header."""
        start = self.fp.tell()
        raw = self.fp.read(4 + 18 + 4 + 4 + 2 * 64)
        h = engine.parse_header(raw)
        if h["layout"] == 6:
            engine.upgrade_legacy_stream(raw)         # raises: unsupported generation
        self.fp.seek(start + h["header_bytes"])
        fs, nCh, numSamples, nMDCTLines = h["sample_rate"], h["n_channels"], h["num_samples"], h["n_mdct_lines"]
        nScaleBits, nMantSizeBits, useSBR, useVQ, nLines = (h["n_scale_bits"], h["n_mant_size_bits"], h["use_sbr"],
                                                            h["use_vq"], h["n_lines"])
        p = CodingParams()
        p.sampleRate, p.nChannels, p.numSamples = fs, nCh, numSamples
        p.nMDCTLines = p.nSamplesPerBlock = nMDCTLines
        p.nScaleBits, p.nMantSizeBits = nScaleBits, nMantSizeBits
        p.sfBands = ScaleFactorBands(nLines)
        p.sfBandsShort = ScaleFactorBands(AssignMDCTLinesFromFreqLimits(128, fs))
        p.useSBR = bool(useSBR)
        p.omittedBands = omitted_bands(p.sfBands) if p.useSBR else []
        p.useVQ = bool(useVQ)
        p.overlapAndAdd = [np.zeros(nMDCTLines, dtype=np.float64) for _ in range(nCh)]
        # the kernels use the band tables they build from (nMDCTLines, sampleRate): refuse any other layout
        engine.context_for_header(dict(sample_rate=fs, n_channels=nCh, use_vq=p.useVQ, use_sbr=p.useSBR,
                                       n_scale_bits=nScaleBits, n_mant_size_bits=nMantSizeBits,
                                       n_lines=tuple(nLines), n_mdct_lines=nMDCTLines))
        return p

    def WriteFileHeader(self, codingParams):
        """(coder/pacfile.py:300-340), including its numSamples rule: padding is added only when
        numSamples is ALREADY a multiple of the block size, then one more block."""
        cp = codingParams
        self.fp.write(self.tag)
        if not cp.numSamples % cp.nMDCTLines:
            cp.numSamples += cp.nMDCTLines - cp.numSamples % cp.nMDCTLines
        cp.numSamples += cp.nMDCTLines
        self.fp.write(pack("<LHLLHHHH", cp.sampleRate, cp.nChannels, cp.numSamples, cp.nMDCTLines, cp.nScaleBits,
                           cp.nMantSizeBits, cp.useSBR, cp.useVQ))
        cp.sfBands = ScaleFactorBands(AssignMDCTLinesFromFreqLimits(cp.nMDCTLines, cp.sampleRate))
        cp.sfBandsShort = ScaleFactorBands(AssignMDCTLinesFromFreqLimits(128, cp.sampleRate))
        cp.omittedBands = omitted_bands(cp.sfBands) if cp.useSBR else []
        self.fp.write(pack("<L", cp.sfBands.nBands))
        self.fp.write(pack("<%dH" % cp.sfBands.nBands, *cp.sfBands.nLines.tolist()))
        cp.priorBlock = [np.zeros(cp.nMDCTLines, dtype=np.float64) for _ in range(cp.nChannels)]

    # ------------------------------------------------------------------ encode
    def WriteDataBlock(self, data, codingParams, lastTrans=False, curTrans=False, nextTrans=False):
        """One block of signed fractions per channel -> length-prefixed coded records
        (coder/pacfile.py:449-610)."""
        cp = codingParams
        C_ = cp.nChannels
        full = np.stack([np.concatenate((np.asarray(cp.priorBlock[c], dtype=np.float64),
                                         np.asarray(data[c], dtype=np.float64))) for c in range(C_)])
        cp.priorBlock = data
        ctx = codec.context_for(cp, cp.useSBR)
        dev = torch.device("cuda", ctx.device)
        x = torch.from_numpy(np.ascontiguousarray(full)).to(dev)
        fl = torch.full((C_,), codec._flag(lastTrans, curTrans, nextTrans), dtype=torch.uint8, device=dev)
        stride = _lib.lib.pac_record_stride()
        rec = torch.zeros((C_, stride), dtype=torch.uint8, device=dev)
        nbytes = torch.zeros(C_, dtype=torch.int32, device=dev)
        zsub = torch.zeros(C_, dtype=torch.uint8, device=dev)
        err = _lib.C.c_int(0)
        rc = _lib.lib.pac_encode_blocks(ctx.h, ptr(x), ptr(fl), C_, ptr(rec), ptr(nbytes), ptr(zsub), _lib.C.byref(err),
                                        _lib.stream_ptr())
        _lib.check_soft(rc, err)
        if cp.useSBR and not curTrans:            # BitAlloc_SBR's in-place edit (coder/bitalloc.py:142-143)
            for b in cp.omittedBands:
                cp.sfBands.nLines[b] = 1
        if bool(zsub.any()):
            return                                # an all-zero short sub-frame drops the block (:503-506, :530-533)
        nb = nbytes.cpu().numpy()
        host = rec.cpu().numpy()
        for c in range(C_):
            self.fp.write(pack("<L", int(nb[c])))
            self.fp.write(host[c, :int(nb[c])].tobytes())

    def Close(self, codingParams):
        """Flushes the last block through the MDCT delay when writing (coder/pacfile.py:612-625)."""
        if self.fp.mode == "wb":
            z = [np.zeros(codingParams.nMDCTLines, dtype=np.float64) for _ in range(codingParams.nChannels)]
            self.WriteDataBlock(z, codingParams)
        self.fp.close()

    # ------------------------------------------------------------------ decode
    def ReadDataBlock(self, codingParams):
        """Next block of every channel as signed fractions, overlap-added with the previous block's
        tail; at the end of file the pending tails, then None (coder/pacfile.py:231-298)."""
        cp = codingParams
        C_ = cp.nChannels
        recs = []
        for _ in range(C_):
            s = self.fp.read(4)
            if not s:
                if cp.overlapAndAdd:
                    ola, cp.overlapAndAdd = cp.overlapAndAdd, 0
                    return ola
                return None
            n = unpack("<L", s)[0]
            buf = self.fp.read(n)
            if len(buf) < n:
                raise RuntimeError("Only read a partial block of coded PACFile data")
            recs.append(buf)
        ctx = codec.context_for(cp, cp.useSBR)
        dev = torch.device("cuda", ctx.device)
        stride = (max(len(r) for r in recs) + 3) // 4 * 4
        host = np.zeros((C_, stride), dtype=np.uint8)
        for c, r in enumerate(recs):
            host[c, :len(r)] = np.frombuffer(r, dtype=np.uint8)
        rec = torch.from_numpy(host).to(dev)
        nb = torch.tensor([len(r) for r in recs], dtype=torch.int32, device=dev)
        out = torch.empty((C_, 2 * cp.nMDCTLines), dtype=torch.float64, device=dev)
        err = _lib.C.c_int(0)
        rc = _lib.lib.pac_decode_blocks(ctx.h, ptr(rec), stride, ptr(nb), C_, ptr(out), _lib.C.byref(err), _lib.stream_ptr())
        _lib.check_soft(rc, err)
        y = out.cpu().numpy()
        H = cp.nMDCTLines
        data = []
        for c in range(C_):
            data.append(np.add(cp.overlapAndAdd[c], y[c, :H]))
            cp.overlapAndAdd[c] = y[c, H:].copy()
        return data


# ---------------------------------------------------------------------- whole files, batched
def _read_wav(path):
    """WAV -> (params, int16 PCM [C, n_read], n_header) the way the reference's driver loop reads it:
    PACFile.WriteFileHeader (coder/pacfile.py:309-315) adds one or two blocks to the SAME
    codingParams.numSamples that bounds PCMFile.ReadDataBlock (coder/pcmfile.py:66-80), so the loop reads
    up to 2048 samples past the data chunk - trailing RIFF chunks are coded as PCM, a short read is zero
    padded.  n_header is the data-chunk count the .pac header is derived from."""
    f = __import__("audio_codec_b200.pcmfile", fromlist=["PCMFile"]).PCMFile(str(path))
    p = f.OpenForReading()
    n = p.numSamples
    n_read = n + 1024 + (1024 if n % 1024 == 0 else 0)
    raw = f.fp.read(n_read * p.nChannels * 2)
    f.fp.close()
    raw = raw + b"\0" * (-len(raw) % (2 * p.nChannels))
    pcm = np.frombuffer(raw, dtype="<i2").reshape(-1, p.nChannels).T
    return p, np.ascontiguousarray(pcm), n


def encode_file(in_wav, out_pac, data_rate, use_vq=True, use_sbr=None):
    """WAV -> .pac with the configuration of the reference's driver (coder/pacfile.py:699-707):
    the whole file in one batched device pass (transient detection included)."""
    p, pcm, n_header = _read_wav(in_wav)
    ctx = _lib.Context(sample_rate=p.sampleRate, n_channels=p.nChannels, rate_kbps=data_rate, use_vq=use_vq, use_sbr=use_sbr)
    buf = engine.encode_streams(ctx, [pcm], hdr_samples=[n_header])[0]
    with open(out_pac, "wb") as fp:
        fp.write(buf)
    return len(buf)


def decode_file(in_pac, out_wav):
    """.pac -> 16-bit WAV; the whole file in one batched device pass."""
    buf = open(in_pac, "rb").read()
    h = engine.parse_header(buf)
    pcm = engine.decode_streams([buf])[0]
    n = min(pcm.shape[1], h["num_samples"]) if h["num_samples"] else pcm.shape[1]
    f = __import__("audio_codec_b200.pcmfile", fromlist=["PCMFile"]).PCMFile(str(out_wav))
    cp = CodingParams()
    cp.nChannels, cp.sampleRate, cp.bitsPerSample, cp.numSamples = h["n_channels"], h["sample_rate"], 16, h["num_samples"]
    f.fp = open(str(out_wav), "wb")
    f.WriteFileHeader(cp)
    f.fp.write(np.ascontiguousarray(pcm.T.astype("<i2")).tobytes())
    f.fp.close()
    return pcm.shape[1]


def encode_decode_test(in_file, data_rate, output_dir="."):
    """The reference's driver (coder/pacfile.py:674-762) on the batched path: encode `in_file` at
    `data_rate` kb/s per channel into <stem>_coded_<rate>.pac and decode it into <stem>_<rate>.wav."""
    import os
    stem = os.path.splitext(os.path.basename(str(in_file)))[0]
    pac = os.path.join(output_dir, f"{stem}_coded_{data_rate}.pac")
    wav = os.path.join(output_dir, f"{stem}_{data_rate}.wav")
    encode_file(in_file, pac, data_rate)
    decode_file(pac, wav)
    return pac, wav
