"""Restatement, over frame records, of what apps/m17-demod.cpp keeps on top of the decoder's callbacks
(TEST INFRASTRUCTURE): packet reassembly with sequence and CRC check (:80-175, :207-253), PRBS9 bit-error
accounting (decode_bert :287-305, Util.h:320-413) and its reset by the diagnostic callback (:353-355).
Pinned against the unmodified reference where it has the code (PRBS9: test_consumers_cpu.py)."""
import numpy as np

import oracle_lib as ol


def crc_x25(data: bytes) -> int:
    reg = 0xFFFF
    for b in data:
        reg ^= b
        for _ in range(8):
            reg = (reg >> 1) ^ 0x8408 if reg & 1 else reg >> 1
    return (~reg) & 0xFFFF


class Packets:
    """current_packet / packet_frame_counter of the CLI, fed one record at a time."""

    def __init__(self):
        self.cur = b""
        self.counter = 0
        self.done = []           # (sample_pos, fcs_ok, bytes) per final segment
        self.seq_errors = 0

    def feed(self, r):
        t = ol.TYPE_NAMES[int(r["type"])]
        if t == "LSF":
            lsf = bytes(r["lsf"])
            self.cur, self.counter = b"", 0
            if not (lsf[13] & 1) and ((lsf[13] >> 1) & 3) != 1:
                self.cur = lsf
        elif t in ("BASIC_PACKET", "FULL_PACKET"):
            seg = bytes(r["payload"][:26])
            number = (seg[25] & 0x7F) >> 2
            if seg[25] & 0x80:
                self.cur += seg[:min(number, 25)]
                self.done.append((int(r["sample_pos"]), crc_x25(self.cur) == 0x0F47, self.cur))
            elif number != self.counter:
                self.seq_errors += 1
            else:
                self.counter += 1
                self.cur += seg[:25]


class Prbs9:
    """Util.h:320-413."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.state, self.synced, self.sync_count = 1, False, 0
        self.bit_count = self.err_count = 0
        self.history = [0] * 128
        self.hist_count = self.hist_pos = 0

    def validate(self, bit):
        if not self.synced:
            r = (bit ^ (self.state >> 8) ^ (self.state >> 4)) & 1
            self.state = ((self.state << 1) | bit) & 0x1FF
            if r:
                self.sync_count = 0
            else:
                self.sync_count += 1
                if self.sync_count == 18:
                    self.synced = True
                    self.bit_count += 18
                    self.history = [0] * 128
                    self.hist_count = self.hist_pos = self.sync_count = 0
            return r
        g = ((self.state >> 8) ^ (self.state >> 4)) & 1
        self.state = ((self.state << 1) | g) & 0x1FF
        err = bit ^ g
        self.bit_count += 1
        self.hist_count -= self.history[self.hist_pos]
        if err:
            self.err_count += 1
            self.hist_count += 1
            self.history[self.hist_pos] = 1
            if self.hist_count >= 25:
                self.synced = False
        else:
            self.history[self.hist_pos] = 0
        self.hist_pos = (self.hist_pos + 1) % 128
        return err

    def frame(self, bert25: bytes):
        for k in range(197):
            self.validate((bert25[k >> 3] >> (7 - (k & 7))) & 1)


def diag_events(trace):
    """(samples, dcd) of every diagnostic callback the reference makes (M17Demodulator.h:675-689,742-752), from the
    oracle's per-sample trace: count_ returns to 0 exactly on the samples that fire one."""
    idx = np.nonzero(trace["count"][1920:] == 0)[0] + 1920
    return [(int(i) + 1, bool(trace["dcd"][i])) for i in idx]


def run_consumers(records, trace):
    """Feed one stream's records and diagnostic callbacks in the reference's order (the frame callback of a sample
    before its diagnostic callback).  Returns (Packets, Prbs9, list of (errors, bits) the CLI would print)."""
    pk, pr, lines = Packets(), Prbs9(), []
    ev = [(int(r["sample_pos"]) + 1, 0, r) for r in records] + [(s, 1, dcd) for s, dcd in diag_events(trace)]
    ev.sort(key=lambda e: (e[0], e[1]))
    for _, kind, x in ev:
        if kind == 0:
            pk.feed(x)
            if ol.TYPE_NAMES[int(x["type"])] == "BERT":
                pr.frame(bytes(x["payload"][:25]))
        else:
            if not x and pr.bit_count > 0:
                pr.reset()
            if pr.bit_count > 0:
                lines.append((pr.err_count, pr.bit_count))
    return pk, pr, lines
