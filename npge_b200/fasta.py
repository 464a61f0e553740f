"""FASTA records for the device ingest.

The reference reads FASTA with FastaReader (src/util/FastaReader.cpp): a record
starts at a line beginning with '>', the first word of that line is the name,
the rest the description; the body runs to the next such line.  Only the record
boundaries are found here (vectorised, on the host); the bodies go to the GPU
as they are -- npge_anchor_upload_raw drops the line breaks and applies
Sequence::to_atgcn there."""
import numpy as np


def split_records(data):
    """bytes / uint8 array of a FASTA file -> (names, descriptions, raw, offsets):
    raw[offsets[i]:offsets[i + 1]] is the body of record i, line breaks included."""
    buf = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(
        data, np.ndarray) else np.ascontiguousarray(data, dtype=np.uint8)
    n = len(buf)
    gt = np.flatnonzero(buf == ord(">"))
    at_line_start = (gt == 0) | (buf[np.maximum(gt, 1) - 1] == ord("\n"))
    heads = gt[at_line_start]
    nl = np.flatnonzero(buf == ord("\n"))
    names, descriptions, begins, ends = [], [], [], []
    for i, h in enumerate(heads):
        j = np.searchsorted(nl, h)
        line_end = int(nl[j]) if j < len(nl) else n
        header = bytes(buf[h + 1:line_end]).decode("ascii", "replace").rstrip("\r")
        parts = header.split(None, 1)
        names.append(parts[0] if parts else "")
        descriptions.append(parts[1] if len(parts) > 1 else "")
        begins.append(min(line_end + 1, n))
        ends.append(int(heads[i + 1]) if i + 1 < len(heads) else n)
    # bodies are contiguous pieces of the file: pass the file itself, with one
    # (begin, end) per record squeezed into monotone offsets by dropping headers
    pieces = [buf[b:e] for b, e in zip(begins, ends)]
    offsets = np.zeros(len(pieces) + 1, dtype=np.uint64)
    if pieces:
        offsets[1:] = np.cumsum([len(p) for p in pieces])
        raw = np.concatenate(pieces) if len(pieces) > 1 else pieces[0].copy()
    else:
        raw = np.zeros(0, dtype=np.uint8)
    return names, descriptions, np.ascontiguousarray(raw), offsets
