"""A board in shared host memory where every worker of one node posts the newest round it has started.

The reference worker starts a round by itself when a peer's ScatterBlock / ReduceBlock for a round beyond its maxRound
arrives (AllreduceWorker.scala:123-126, AllreduceSpec.scala:106-124): progress of the peers reaches a worker through the
data path, not only through the master.  On the device plane the data path is NVLink stores into this worker's ring -- no
host sees them -- so the same information travels next to it: one 8-byte word per worker in a file under /dev/shm, written
by the worker's host thread when it launches a round, read by its peers' event loops.  One NVSwitch domain is one host, so
shared memory reaches every peer a worker can have.
"""
from __future__ import annotations

import mmap
import os
import struct


class PeerBoard:
    def __init__(self, session: str, world: int, rank: int, initial: int = -1, directory: str = "/dev/shm"):
        self.world, self.rank = world, rank
        safe = "".join(c if c.isalnum() or c in "-_." else "_" for c in session)
        self.path = os.path.join(directory if os.path.isdir(directory) else "/tmp", f"aar_board_{os.getuid()}_{safe}")
        size = 8 * world
        fd = os.open(self.path, os.O_RDWR | os.O_CREAT, 0o600)
        try:
            if os.fstat(fd).st_size < size:
                os.ftruncate(fd, size)   # new bytes read as zero = round -1 (stored value is round + 1)
            self.mm = mmap.mmap(fd, size)
        finally:
            os.close(fd)
        # A fresh incarnation: whatever an earlier job with the same session name left in this worker's word is stale (a stale
        # word ahead of the master's clock would make every peer self-start rounds nobody asked for).  Callers create the board
        # before the barrier that precedes the first launch, so every word has been reset before anybody reads one.
        struct.pack_into("<q", self.mm, 8 * self.rank, initial + 1)

    def post(self, newest_round: int) -> None:
        """This worker has started every round up to `newest_round` (monotone)."""
        off = 8 * self.rank
        cur = struct.unpack_from("<q", self.mm, off)[0]
        if newest_round + 1 > cur:
            struct.pack_into("<q", self.mm, off, newest_round + 1)

    def newest(self) -> int:
        """The newest round any worker on the board has started (-1: none)."""
        return max(struct.unpack_from(f"<{self.world}q", self.mm, 0)) - 1

    def close(self, unlink: bool = False) -> None:
        try:
            self.mm.close()
        except (BufferError, ValueError):
            pass
        if unlink:
            try:
                os.unlink(self.path)
            except OSError:
                pass
