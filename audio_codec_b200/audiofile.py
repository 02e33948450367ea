"""Drop-in for the reference's `audiofile` module (coder/audiofile.py:51-93): the
CodingParams attribute bag and the AudioFile base class."""

__all__ = ["CodingParams", "AudioFile"]


class CodingParams:
    """Attribute bag filled by the file objects (coder/audiofile.py:51-53)."""
    pass


class AudioFile:
    """Open / read / write / close skeleton the PCM and PAC file classes share
    (coder/audiofile.py:56-93)."""

    def __init__(self, filename):
        self.filename = filename

    def OpenForReading(self):
        self.fp = open(self.filename, "rb")
        return self.ReadFileHeader()

    def OpenForWriting(self, codingParams):
        self.fp = open(self.filename, "wb")
        self.WriteFileHeader(codingParams)

    def Close(self, codingParams):
        self.fp.close()

    def ReadFileHeader(self):
        return CodingParams()

    def WriteFileHeader(self, codingParams):
        pass

    def ReadDataBlock(self, codingParams):
        pass

    def WriteDataBlock(self, data, codingParams):
        pass
