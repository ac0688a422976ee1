"""Host-side exchange between the ranks of a dist=True run, without any framework: a star of plain TCP
sockets on the loopback / MASTER_ADDR interface.

The reference hands this to tf.distribute.MirroredStrategy (models.py:219, fit.py:149-153), which lives in
one process.  Here every GPU has its own process (torchrun-style RANK / WORLD_SIZE / MASTER_ADDR /
MASTER_PORT environment) and the host side has exactly two things to move between them, both tiny and both
outside the training loop: rank 0's 128-byte NCCL unique id plus its initial weights (broadcast), and the
ranks' shards of a self-adaptive lambda vector when the user reads `model.lambdas` back (all-gather).
Everything per step goes through NCCL inside libtdq_b200.

Rank 0 listens on TDQ_RDV_PORT (default MASTER_PORT + 29, so that it does not collide with the store a
launcher such as torchrun keeps on MASTER_PORT itself); the other ranks connect with retries.
"""
import os
import socket
import struct
import time

_MAGIC = b"TDQRDV1\0"


def _send(sock, payload):
    sock.sendall(struct.pack("<Q", len(payload)) + payload)


def _recv_exact(sock, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(1 << 20, n - len(buf)))
        if not chunk:
            raise ConnectionError("rendezvous peer closed the connection")
        buf += chunk
    return bytes(buf)


def _recv(sock):
    (n,) = struct.unpack("<Q", _recv_exact(sock, 8))
    return _recv_exact(sock, n)


class Rendezvous:
    def __init__(self, rank, world, addr=None, port=None, timeout=300.0):
        self.rank, self.world = rank, world
        self.peers = {}
        self.sock = None
        if world == 1:
            return
        addr = addr or os.environ.get("MASTER_ADDR", "127.0.0.1")
        port = int(port or os.environ.get("TDQ_RDV_PORT", int(os.environ.get("MASTER_PORT", "29500")) + 29))
        deadline = time.time() + timeout
        if rank == 0:
            srv = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            srv.bind((addr, port))
            srv.listen(world)
            srv.settimeout(timeout)
            while len(self.peers) < world - 1:
                c, _ = srv.accept()
                c.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                hello = _recv_exact(c, len(_MAGIC) + 4)
                if hello[:len(_MAGIC)] != _MAGIC:
                    c.close()
                    continue
                (r,) = struct.unpack("<i", hello[len(_MAGIC):])
                self.peers[r] = c
            srv.close()
        else:
            while True:
                try:
                    s = socket.create_connection((addr, port), timeout=5.0)
                    s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                    s.sendall(_MAGIC + struct.pack("<i", rank))
                    s.settimeout(timeout)
                    self.sock = s
                    break
                except OSError:
                    if time.time() > deadline:
                        raise TimeoutError("rendezvous: rank 0 is not listening on %s:%d" % (addr, port))
                    time.sleep(0.05)

    def broadcast(self, payload):
        """rank 0's bytes to every rank"""
        if self.world == 1:
            return payload
        if self.rank == 0:
            for r in sorted(self.peers):
                _send(self.peers[r], payload)
            return payload
        return _recv(self.sock)

    def all_gather(self, payload):
        """every rank's bytes, in rank order, on every rank"""
        if self.world == 1:
            return [payload]
        if self.rank == 0:
            parts = [payload] + [_recv(self.peers[r]) for r in sorted(self.peers)]
            blob = struct.pack("<I", len(parts)) + b"".join(struct.pack("<Q", len(p)) + p for p in parts)
            for r in sorted(self.peers):
                _send(self.peers[r], blob)
            return parts
        _send(self.sock, payload)
        blob = _recv(self.sock)
        (k,) = struct.unpack("<I", blob[:4])
        parts, off = [], 4
        for _ in range(k):
            (n,) = struct.unpack("<Q", blob[off:off + 8])
            parts.append(blob[off + 8:off + 8 + n])
            off += 8 + n
        return parts

    def barrier(self):
        self.all_gather(b"")

    def close(self):
        for c in self.peers.values():
            c.close()
        if self.sock:
            self.sock.close()
        self.peers, self.sock = {}, None
