"""Symmetric (peer-mapped) HBM regions across one-process-per-GPU ranks, plus the optional NVLS multicast view.

What Akka remoting gives the reference -- an address for every worker that every other worker can send to
(application.conf:5-16, AllreduceMaster.scala:70-79) -- is here a CUDA VMM allocation per rank whose POSIX file
descriptor is passed to the peers (SCM_RIGHTS over a Unix-domain socket) and mapped into their address space.
After `SymmetricMemory(...)` every rank holds `ptrs[j]`, rank j's region as seen from this GPU over NVLink.

The rendezvous is a tiny star over an abstract Unix socket owned by rank 0; it needs neither NCCL nor
torch.distributed (the framework's own master/worker CLIs use it as is).
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import struct
import time

from ..ops import lib as L


class UnixRendezvous:
    """Star rendezvous on one host: fd exchange, small blob all-gather and barrier."""

    def __init__(self, rank: int, world: int, session: str | None = None, timeout: float = 120.0):
        self.rank, self.world = rank, world
        session = session or os.environ.get("AAR_SESSION") or f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
        self.addr = "\0aar_rdv_" + session
        self.conns: list[socket.socket] = []
        self.conn: socket.socket | None = None
        if world == 1:
            return
        if rank == 0:
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(self.addr)
            srv.listen(world)
            srv.settimeout(timeout)
            by_rank = {}
            while len(by_rank) < world - 1:
                c, _ = srv.accept()
                c.settimeout(timeout)
                # The abstract socket namespace has no file permissions, and whoever completes this rendezvous receives the
                # file descriptors of this job's device memory: only a process of the same user (or root) may take part, and
                # only with a rank that exists and has not been claimed.
                try:
                    _pid, uid, _gid = struct.unpack("3i", c.getsockopt(socket.SOL_SOCKET, socket.SO_PEERCRED, struct.calcsize("3i")))
                except OSError:
                    uid = None
                if uid is None or uid not in (os.getuid(), 0):
                    c.close()
                    continue
                try:
                    (r,) = struct.unpack("i", self._recvn(c, 4))
                except (ConnectionError, socket.timeout, OSError):
                    c.close()
                    continue
                if not (1 <= r < world) or r in by_rank:
                    c.close()   # out-of-range or duplicate rank: not one of ours
                    continue
                by_rank[r] = c
            self.conns = [by_rank[r] for r in range(1, world)]
            srv.close()
        else:
            deadline = time.time() + timeout
            while True:
                s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                try:
                    s.connect(self.addr)
                    break
                except (FileNotFoundError, ConnectionRefusedError):
                    s.close()
                    if time.time() > deadline:
                        raise TimeoutError(f"rendezvous: rank 0 never opened {self.addr!r}")
                    time.sleep(0.05)
            s.settimeout(timeout)
            s.sendall(struct.pack("i", rank))
            self.conn = s

    @staticmethod
    def _recvn(s: socket.socket, n: int) -> bytes:
        buf = b""
        while len(buf) < n:
            part = s.recv(n - len(buf))
            if not part:
                raise ConnectionError("rendezvous peer closed the connection")
            buf += part
        return buf

    def barrier(self) -> None:
        if self.world == 1:
            return
        if self.rank == 0:
            for c in self.conns:
                self._recvn(c, 1)
            for c in self.conns:
                c.sendall(b"b")
        else:
            self.conn.sendall(b"b")
            self._recvn(self.conn, 1)

    def all_gather_bytes(self, blob: bytes) -> list[bytes]:
        if self.world == 1:
            return [blob]
        if self.rank == 0:
            table = [blob]
            for c in self.conns:
                (n,) = struct.unpack("q", self._recvn(c, 8))
                table.append(self._recvn(c, n))
            packed = b"".join(struct.pack("q", len(b)) + b for b in table)
            for c in self.conns:
                c.sendall(packed)
            return table
        self.conn.sendall(struct.pack("q", len(blob)) + blob)
        out = []
        for _ in range(self.world):
            (n,) = struct.unpack("q", self._recvn(self.conn, 8))
            out.append(self._recvn(self.conn, n))
        return out

    def all_gather_fds(self, fd: int) -> list[int]:
        """Every rank contributes one fd and receives a duplicate of every rank's fd (its own slot keeps `fd`)."""
        if self.world == 1:
            return [fd]
        if self.rank == 0:
            table = [fd]
            for c in self.conns:
                _, fds, _, _ = socket.recv_fds(c, 1, 1)
                table.append(fds[0])
                c.sendall(b"a")
            for c in self.conns:
                for r in range(self.world):
                    socket.send_fds(c, [b"f"], [table[r]])
                    self._recvn(c, 1)  # lock-step: one SCM_RIGHTS message in flight at a time
            return table
        socket.send_fds(self.conn, [b"f"], [fd])
        self._recvn(self.conn, 1)
        out = []
        for r in range(self.world):
            _, fds, _, _ = socket.recv_fds(self.conn, 1, 1)
            self.conn.sendall(b"a")
            out.append(fds[0])
        os.close(out[self.rank])
        out[self.rank] = fd
        return out

    def broadcast_fd(self, fd: int | None) -> int:
        """Rank 0's fd reaches every rank."""
        if self.world == 1:
            return fd
        if self.rank == 0:
            for c in self.conns:
                socket.send_fds(c, [b"f"], [fd])
                self._recvn(c, 1)
            return fd
        _, fds, _, _ = socket.recv_fds(self.conn, 1, 1)
        self.conn.sendall(b"a")
        return fds[0]

    def close(self) -> None:
        for c in self.conns:
            c.close()
        if self.conn:
            self.conn.close()


class SymmetricMemory:
    """One VMM allocation per rank, mapped by all ranks; optionally bound to a multicast object (NVLS)."""

    def __init__(self, rdv: UnixRendezvous, nbytes: int, device: int, want_multicast: bool = True):
        self.lib = L.load()
        self.rdv = rdv
        self.rank, self.world, self.device = rdv.rank, rdv.world, device
        lib = self.lib
        mc_ok = bool(want_multicast and self.world > 1 and lib.aar_mc_supported(device))
        # every rank must agree, otherwise nobody uses multicast
        flags = rdv.all_gather_bytes(b"\1" if mc_ok else b"\0")
        mc_ok = all(f == b"\1" for f in flags)
        gran = C.c_size_t(0)
        L.check(lib.aar_vmm_granularity(device, int(mc_ok), self.world, C.byref(gran)), "aar_vmm_granularity")
        self.nbytes = (nbytes + gran.value - 1) // gran.value * gran.value
        self.local = L.AarVmm()
        L.check(lib.aar_vmm_create(device, self.nbytes, C.byref(self.local)), "aar_vmm_create")
        fds = rdv.all_gather_fds(self.local.fd)
        self.imports: list[L.AarVmm | None] = [None] * self.world
        self.ptrs: list[int] = [0] * self.world
        for r in range(self.world):
            if r == self.rank:
                self.ptrs[r] = self.local.ptr
                continue
            v = L.AarVmm()
            L.check(lib.aar_vmm_import(device, fds[r], self.nbytes, C.byref(v)), f"aar_vmm_import(rank {r})")
            os.close(fds[r])
            self.imports[r] = v
            self.ptrs[r] = v.ptr
        # Load the kernel module now (the first launch of a process pays for lazy loading; ranks that differ by seconds
        # there would look like stragglers to each other) and leave the constructor together.
        import torch

        with torch.cuda.device(device):
            L.check(lib.aar_launch_delay(1, None), "aar_launch_delay")
            torch.cuda.synchronize(device)
        rdv.barrier()
        self.mc_ptr = 0
        self.mc_handle = 0
        self.mc_error = ""
        if mc_ok:
            self._setup_multicast()

    def _agree(self, ok: bool) -> bool:
        """Collective AND: every step of the multicast set-up either succeeds on all ranks or is abandoned by all."""
        return all(o == b"\1" for o in self.rdv.all_gather_bytes(b"\1" if ok else b"\0"))

    def _setup_multicast(self) -> None:
        """NVLS set-up; the capability probe may say yes and the host still refuse (containers, fabric state): any
        failure leaves mc_ptr = 0 on every rank and the P2P paths are used."""
        lib, rdv = self.lib, self.rdv
        handle = C.c_uint64(0)
        fd = -1
        ok = True
        if self.rank == 0:
            f = C.c_int(-1)
            try:
                L.check(lib.aar_mc_create(self.world, self.nbytes, C.byref(handle), C.byref(f)), "aar_mc_create")
                fd = f.value
            except L.AarError as e:
                self.mc_error, ok = str(e), False
        if not self._agree(ok):
            return
        fd = rdv.broadcast_fd(fd)
        try:
            if self.rank != 0:
                L.check(lib.aar_mc_import(fd, C.byref(handle)), "aar_mc_import")
            L.check(lib.aar_mc_add_device(handle.value, self.device), "aar_mc_add_device")
        except L.AarError as e:
            self.mc_error, ok = str(e), False
        os.close(fd)
        self.mc_handle = handle.value
        if not self._agree(ok):  # doubles as the barrier: every device is added before anyone binds
            return
        ptr = C.c_void_p(0)
        try:
            L.check(lib.aar_mc_bind_and_map(self.mc_handle, self.device, C.byref(self.local), self.nbytes, C.byref(ptr)), "aar_mc_bind_and_map")
        except L.AarError as e:
            self.mc_error, ok = str(e), False
        if self._agree(ok):
            self.mc_ptr = ptr.value or 0

    @property
    def has_multicast(self) -> bool:
        return bool(self.mc_ptr)

    def tensor(self, rank: int, offset: int, nbytes: int, dtype):
        """A torch view of (part of) a rank's region as mapped here (debug / checkpoint)."""
        import torch

        class _Holder:  # __cuda_array_interface__ v3
            pass

        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (self.ptrs[rank] + offset, False), "version": 3}
        return torch.as_tensor(h, device=f"cuda:{self.device}").view(dtype)

    def close(self) -> None:
        lib = self.lib
        if self.mc_handle:
            lib.aar_mc_release(self.mc_handle, C.c_void_p(self.mc_ptr), self.nbytes)
            self.mc_handle = 0
        for v in self.imports:
            if v is not None:
                lib.aar_vmm_release(C.byref(v))
        lib.aar_vmm_release(C.byref(self.local))
