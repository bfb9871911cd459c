"""``mpi4py.MPI`` look-alike for ONE node, used only to run the unmodified reference (``baseline/_ref``).

The image has neither an MPI library nor mpi4py, so the reference's ``from mpi4py import MPI`` is answered by
this module. It implements the subset the reference calls (pytorch_code/sync_replicas_master_nn.py:131-166,
211-303; distributed_worker.py:156-236; model_ops/lenet.py:106-190; model_ops/resnet_split.py:235-440):

* ``COMM_WORLD.Get_rank / Get_size``
* pickled-object messages ``send / recv / isend / irecv`` (lower case), request ``wait()``
* buffer messages ``Send / Recv / Isend / Irecv`` with ``[ndarray, MPI.DOUBLE]`` specs, ``Bcast``, ``Barrier``
* ``Request.Wait / Test / Cancel``, ``Request.Waitany(requests, status)``, ``Status``, ``Iprobe``
* ``ANY_SOURCE``, ``ANY_TAG``, ``UNDEFINED`` and the datatype handles.

Semantics kept from MPI: matching on (source, tag) with wild cards, posted receives matched in post order,
non-overtaking delivery per (source, destination) pair, receives cancellable while unmatched, sends that
complete asynchronously (a sender thread per destination plays the role of MPI's progress engine, so an
``Isend`` of a large gradient overlaps with the caller's next backward layer as it would with mpirun).

Transport: one AF_UNIX stream socket per ordered pair of ranks under ``$PDNN_MPI_DIR``; the payload of a buffer
message is received straight into the posted buffer when a receive is already waiting (one copy, like the
shared-memory path of an MPI library), otherwise into a temporary that is copied on match.

Rank / size come from ``PDNN_MPI_RANK`` / ``PDNN_MPI_SIZE`` (set by ``baseline/run_reference.py``); without
them the module behaves as a one-rank world so that plain imports work.
"""
import atexit
import os
import pickle
import queue
import socket
import struct
import threading
import time

import numpy as _np

ANY_SOURCE = -2
ANY_TAG = -1
UNDEFINED = -32766
PROC_NULL = -3
SUCCESS = 0

_HDR = struct.Struct("<iiBBq")          # source, tag, context, kind, payload bytes
_K_OBJ, _K_BUF, _K_FIN = 0, 1, 2
_CTX_P2P, _CTX_COLL = 0, 1


class PeerFinished(BaseException):
    """Raised out of a blocking call after another rank announced the end of the job (``COMM_WORLD.finish()``).
    Derives from BaseException so that the reference's own ``except Exception`` blocks cannot swallow it."""


class Datatype(object):
    def __init__(self, np_dtype, name):
        self.np_dtype = _np.dtype(np_dtype)
        self.name = name

    def Get_size(self):
        return self.np_dtype.itemsize

    def __repr__(self):
        return "<MPI.%s>" % self.name


DOUBLE = Datatype(_np.float64, "DOUBLE")
FLOAT = Datatype(_np.float32, "FLOAT")
INT = Datatype(_np.int32, "INT")
LONG = Datatype(_np.int64, "LONG")
BYTE = Datatype(_np.uint8, "BYTE")
CHAR = Datatype(_np.int8, "CHAR")


class Status(object):
    def __init__(self):
        self.source = ANY_SOURCE
        self.tag = ANY_TAG
        self.error = SUCCESS
        self.count = 0
        self.cancelled = False

    def Get_source(self):
        return self.source

    def Get_tag(self):
        return self.tag

    def Get_error(self):
        return self.error

    def Get_count(self, datatype=BYTE):
        return self.count // datatype.Get_size()

    def Is_cancelled(self):
        return self.cancelled

    def _fill(self, req):
        self.source, self.tag, self.count, self.cancelled = req.src_seen, req.tag_seen, req.count, req.cancelled


def _byte_view(spec, writable):
    """``[array, MPI.DOUBLE]`` / ``[array, count, MPI.DOUBLE]`` / bare array -> flat byte memoryview."""
    arr = spec[0] if isinstance(spec, (list, tuple)) else spec
    if not isinstance(arr, _np.ndarray):
        arr = _np.asarray(arr)
    if not arr.flags["C_CONTIGUOUS"]:
        raise ValueError("mpi4py shim: buffers must be C-contiguous ndarrays")
    if writable and not arr.flags["WRITEABLE"]:
        raise ValueError("mpi4py shim: receive buffer is read-only")
    return memoryview(arr.reshape(-1).view(_np.uint8)), arr


class Request(object):
    """One outstanding send or receive. ``active`` goes False once Wait/Waitany has reported it (MPI sets the
    handle to REQUEST_NULL at that point), so a second Waitany never returns the same request again."""

    def __init__(self, eng, kind, is_recv, src=ANY_SOURCE, tag=ANY_TAG, ctx=_CTX_P2P, view=None, keep=None):
        self.eng, self.kind, self.is_recv = eng, kind, is_recv
        self.src, self.tag, self.ctx = src, tag, ctx
        self.view, self.keep = view, keep
        self.done = False
        self.active = True
        self.cancelled = False
        self.busy = False                  # a reader thread is filling the buffer right now
        self.obj = None
        self.src_seen, self.tag_seen, self.count = ANY_SOURCE, ANY_TAG, 0
        self.error = None

    # -- completion ------------------------------------------------------------------------------------
    def _wait(self, status=None):
        eng = self.eng
        with eng.cv:
            while not self.done:
                eng._check_finished()
                eng.cv.wait(0.2)
            self.active = False
        if self.error is not None:
            raise RuntimeError(self.error)
        if status is not None:
            status._fill(self)
        return self.obj

    def wait(self, status=None):
        return self._wait(status)

    def Wait(self, status=None):
        self._wait(status)
        return True

    def test(self, status=None):
        with self.eng.cv:
            if not self.done:
                return (False, None)
            self.active = False
        if status is not None:
            status._fill(self)
        return (True, self.obj)

    def Test(self, status=None):
        return self.test(status)[0]

    def Cancel(self):
        """Receives that are still unmatched are withdrawn; anything already matched (or any send) completes
        normally, which is what MPI allows an implementation to do."""
        eng = self.eng
        with eng.cv:
            if self.is_recv and not self.done and not self.busy and self in eng.posted:
                eng.posted.remove(self)
                self.cancelled = True
                self.done = True
                eng.cv.notify_all()

    cancel = Cancel

    def Free(self):
        self.active = False

    # -- MPI_Waitany / Waitall / Testany ---------------------------------------------------------------
    @staticmethod
    def Waitany(requests, status=None):
        live = [r for r in requests if r is not None]
        if not live:
            return UNDEFINED
        eng = live[0].eng
        with eng.cv:
            while True:
                any_active = False
                for i, r in enumerate(requests):
                    if r is None or not r.active:
                        continue
                    any_active = True
                    if r.done and not r.cancelled:
                        r.active = False
                        if r.error is not None:
                            raise RuntimeError(r.error)
                        if status is not None:
                            status._fill(r)
                        return i
                    if r.done and r.cancelled:
                        r.active = False
                if not any_active:
                    return UNDEFINED
                eng._check_finished()
                eng.cv.wait(0.2)

    waitany = Waitany

    @staticmethod
    def Waitall(requests, statuses=None):
        for r in requests:
            if r is not None:
                r._wait()
        return True

    waitall = Waitall

    @staticmethod
    def Testany(requests, status=None):
        for i, r in enumerate(requests):
            if r is not None and r.active and r.done and not r.cancelled:
                r.active = False
                if status is not None:
                    status._fill(r)
                return (i, True)
        return (UNDEFINED, False)


class _Msg(object):
    __slots__ = ("src", "tag", "ctx", "kind", "data")

    def __init__(self, src, tag, ctx, kind, data):
        self.src, self.tag, self.ctx, self.kind, self.data = src, tag, ctx, kind, data


def _recv_exact(sock, view):
    got, n = 0, len(view)
    while got < n:
        r = sock.recv_into(view[got:], min(n - got, 1 << 24))
        if r == 0:
            return False
        got += r
    return True


class _Sender(threading.Thread):
    """Progress thread for everything this rank sends to one destination (keeps per-pair order)."""

    def __init__(self, eng, dest):
        threading.Thread.__init__(self, daemon=True, name="mpi-shim-send-%d" % dest)
        self.eng, self.dest = eng, dest
        self.q = queue.Queue()
        self.sock = None
        self.start()

    def _connect(self):
        path = self.eng._sock_path(self.dest)
        deadline = time.time() + float(os.environ.get("PDNN_MPI_CONNECT_TIMEOUT", "300"))
        while True:
            try:
                s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                s.setsockopt(socket.SOL_SOCKET, socket.SO_SNDBUF, 8 << 20)
                s.connect(path)
                return s
            except OSError:
                s.close()
                if time.time() > deadline:
                    raise
                time.sleep(0.05)

    def run(self):
        while True:
            item = self.q.get()
            if item is None:
                return
            hdr, payload, req = item
            try:
                if self.sock is None:
                    self.sock = self._connect()
                self.sock.sendall(hdr)
                if payload is not None and len(payload):
                    self.sock.sendall(payload)
            except OSError as e:                              # peer went away: only an error if nobody said goodbye
                if req is not None and not self.eng.finished:
                    req.error = "mpi4py shim: send to rank %d failed: %s" % (self.dest, e)
            if req is not None:
                with self.eng.cv:
                    req.done = True
                    req.keep = None
                    self.eng.cv.notify_all()


class _Engine(object):
    def __init__(self, rank, size, rdv):
        self.rank, self.size, self.rdv = rank, size, rdv
        self.cv = threading.Condition()
        self.posted, self.unexpected = [], []
        self.finished = False
        self.senders = {}
        self.coll_seq = 0
        self.bytes_sent = 0
        self.bytes_recv = 0
        self.listener = None
        if size > 1:
            path = self._sock_path(rank)
            try:
                os.unlink(path)
            except OSError:
                pass
            self.listener = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            self.listener.bind(path)
            self.listener.listen(size + 4)
            threading.Thread(target=self._accept_loop, daemon=True, name="mpi-shim-accept").start()

    def _sock_path(self, r):
        return os.path.join(self.rdv, "rank%d.sock" % r)

    def _check_finished(self):
        if self.finished:
            raise PeerFinished("another rank ended the job")

    # -- receive side ------------------------------------------------------------------------------------
    def _accept_loop(self):
        while True:
            try:
                conn, _ = self.listener.accept()
            except OSError:
                return
            conn.setsockopt(socket.SOL_SOCKET, socket.SO_RCVBUF, 8 << 20)
            threading.Thread(target=self._reader, args=(conn,), daemon=True, name="mpi-shim-recv").start()

    @staticmethod
    def _matches(req, src, tag, ctx):
        return req.ctx == ctx and (req.src == ANY_SOURCE or req.src == src) and (req.tag == ANY_TAG or req.tag == tag)

    def _take_posted(self, src, tag, ctx):
        for i, r in enumerate(self.posted):
            if self._matches(r, src, tag, ctx):
                del self.posted[i]
                r.busy = True
                return r
        return None

    def _deliver(self, req, src, tag, kind, data):
        """cv held. ``data`` is None when the payload was already received into req.view."""
        n = 0
        if data is not None:
            n = len(data)
            if kind == _K_OBJ or req.view is None:
                req.obj = pickle.loads(bytes(data)) if kind == _K_OBJ else bytes(data)
            elif n > len(req.view):
                req.error = "mpi4py shim: message of %d bytes truncated by a %d-byte receive" % (n, len(req.view))
            else:
                req.view[:n] = data
        req.src_seen, req.tag_seen = src, tag
        if data is not None:
            req.count = n
        req.busy = False
        req.done = True
        self.cv.notify_all()

    def _reader(self, conn):
        hdr = bytearray(_HDR.size)
        hv = memoryview(hdr)
        while True:
            if not _recv_exact(conn, hv):
                return
            src, tag, ctx, kind, n = _HDR.unpack(hdr)
            if kind == _K_FIN:
                with self.cv:
                    self.finished = True
                    self.cv.notify_all()
                continue
            with self.cv:
                req = self._take_posted(src, tag, ctx)
            if req is not None and kind == _K_BUF and req.view is not None and n <= len(req.view):
                ok = _recv_exact(conn, req.view[:n]) if n else True    # straight into the user's buffer
                with self.cv:
                    req.count = n
                    self.bytes_recv += n
                    if not ok:
                        req.error = "mpi4py shim: connection closed inside a message"
                    self._deliver(req, src, tag, kind, None)
                if not ok:
                    return
                continue
            data = bytearray(n)
            if n and not _recv_exact(conn, memoryview(data)):
                return
            with self.cv:
                self.bytes_recv += n
                if req is None:
                    req = self._take_posted(src, tag, ctx)     # one may have been posted while we were reading
                if req is not None:
                    self._deliver(req, src, tag, kind, data)
                else:
                    self.unexpected.append(_Msg(src, tag, ctx, kind, data))
                    self.cv.notify_all()

    def post_recv(self, req):
        with self.cv:
            for i, m in enumerate(self.unexpected):
                if self._matches(req, m.src, m.tag, m.ctx):
                    del self.unexpected[i]
                    self._deliver(req, m.src, m.tag, m.kind, m.data)
                    return req
            self.posted.append(req)
        return req

    def probe(self, src, tag, ctx, status):
        with self.cv:
            for m in self.unexpected:
                if m.ctx == ctx and (src == ANY_SOURCE or src == m.src) and (tag == ANY_TAG or tag == m.tag):
                    if status is not None:
                        status.source, status.tag, status.count = m.src, m.tag, len(m.data)
                    return True
        return False

    # -- send side ---------------------------------------------------------------------------------------
    def send(self, dest, tag, ctx, kind, payload, keep=None):
        if dest == self.rank:                                   # self-send: deliver locally
            req = Request(self, kind, False)
            data = bytearray(payload) if payload is not None else bytearray()
            with self.cv:
                r = self._take_posted(self.rank, tag, ctx)
                if r is not None:
                    self._deliver(r, self.rank, tag, kind, data)
                else:
                    self.unexpected.append(_Msg(self.rank, tag, ctx, kind, data))
                req.done = True
                self.cv.notify_all()
            return req
        if not 0 <= dest < self.size:
            raise ValueError("mpi4py shim: invalid destination rank %r" % (dest,))
        n = len(payload) if payload is not None else 0
        req = Request(self, kind, False, keep=keep)
        s = self.senders.get(dest)
        if s is None:
            s = self.senders[dest] = _Sender(self, dest)
        self.bytes_sent += n
        s.q.put((_HDR.pack(self.rank, tag, ctx, kind, n), payload, req))
        return req

    def finish_all(self):
        """Tell every other rank that the job is over (their blocking calls raise PeerFinished)."""
        self.finished = True
        pend = []
        for d in range(self.size):
            if d == self.rank:
                continue
            s = self.senders.get(d)
            if s is None:
                s = self.senders[d] = _Sender(self, d)
            r = Request(self, _K_FIN, False)
            s.q.put((_HDR.pack(self.rank, 0, _CTX_P2P, _K_FIN, 0), None, r))
            pend.append(r)
        t_end = time.time() + 10
        for r in pend:
            while not r.done and time.time() < t_end:
                time.sleep(0.01)

    def close(self):
        if self.listener is not None:
            try:
                self.listener.close()
                os.unlink(self._sock_path(self.rank))
            except OSError:
                pass


class Intracomm(object):
    def __init__(self, eng):
        self._eng = eng
        self._send_hooks = []

    # launcher-side instrumentation: hook(kind, payload, dest, tag) runs before every point-to-point send
    def add_send_hook(self, fn):
        self._send_hooks.append(fn)

    def finish(self):
        self._eng.finish_all()

    def traffic(self):
        return {"bytes_sent": self._eng.bytes_sent, "bytes_recv": self._eng.bytes_recv}

    def Get_rank(self):
        return self._eng.rank

    def Get_size(self):
        return self._eng.size

    rank = property(Get_rank)
    size = property(Get_size)

    def _hooks(self, kind, payload, dest, tag):
        for h in self._send_hooks:
            h(kind, payload, dest, tag)

    # -- pickled objects ---------------------------------------------------------------------------------
    def isend(self, obj, dest, tag=0):
        self._hooks("obj", obj, dest, tag)
        return self._eng.send(dest, tag, _CTX_P2P, _K_OBJ, pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))

    def send(self, obj, dest, tag=0):
        self.isend(obj, dest, tag)._wait()

    def irecv(self, buf=None, source=ANY_SOURCE, tag=ANY_TAG):
        return self._eng.post_recv(Request(self._eng, _K_OBJ, True, source, tag))

    def recv(self, buf=None, source=ANY_SOURCE, tag=ANY_TAG, status=None):
        return self.irecv(None, source, tag)._wait(status)

    # -- buffers -----------------------------------------------------------------------------------------
    def Isend(self, buf, dest, tag=0):
        view, arr = _byte_view(buf, False)
        self._hooks("buf", arr, dest, tag)
        return self._eng.send(dest, tag, _CTX_P2P, _K_BUF, view, keep=arr)

    def Send(self, buf, dest, tag=0):
        self.Isend(buf, dest, tag)._wait()

    def Irecv(self, buf, source=ANY_SOURCE, tag=ANY_TAG):
        view, arr = _byte_view(buf, True)
        return self._eng.post_recv(Request(self._eng, _K_BUF, True, source, tag, view=view, keep=arr))

    def Recv(self, buf, source=ANY_SOURCE, tag=ANY_TAG, status=None):
        self.Irecv(buf, source, tag)._wait(status)

    def Iprobe(self, source=ANY_SOURCE, tag=ANY_TAG, status=None):
        return self._eng.probe(source, tag, _CTX_P2P, status)

    iprobe = Iprobe

    # -- collectives (linear: the root talks to every rank, each on its own sender thread) ----------------
    def _next_coll(self):
        self._eng.coll_seq += 1
        return self._eng.coll_seq

    def Bcast(self, buf, root=0):
        seq = self._next_coll()
        eng = self._eng
        if eng.size == 1:
            return
        if eng.rank == root:
            view, arr = _byte_view(buf, False)
            reqs = [eng.send(d, seq, _CTX_COLL, _K_BUF, view, keep=arr) for d in range(eng.size) if d != root]
            for r in reqs:
                r._wait()
        else:
            view, arr = _byte_view(buf, True)
            eng.post_recv(Request(eng, _K_BUF, True, root, seq, _CTX_COLL, view=view, keep=arr))._wait()

    def bcast(self, obj=None, root=0):
        seq = self._next_coll()
        eng = self._eng
        if eng.size == 1:
            return obj
        if eng.rank == root:
            data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
            reqs = [eng.send(d, seq, _CTX_COLL, _K_OBJ, data) for d in range(eng.size) if d != root]
            for r in reqs:
                r._wait()
            return obj
        return eng.post_recv(Request(eng, _K_OBJ, True, root, seq, _CTX_COLL))._wait()

    def Barrier(self):
        seq = self._next_coll()
        eng = self._eng
        if eng.size == 1:
            return
        if eng.rank == 0:
            for _ in range(eng.size - 1):
                eng.post_recv(Request(eng, _K_OBJ, True, ANY_SOURCE, seq, _CTX_COLL))._wait()
            reqs = [eng.send(d, seq, _CTX_COLL, _K_OBJ, pickle.dumps(None)) for d in range(1, eng.size)]
            for r in reqs:
                r._wait()
        else:
            eng.send(0, seq, _CTX_COLL, _K_OBJ, pickle.dumps(None))._wait()
            eng.post_recv(Request(eng, _K_OBJ, True, 0, seq, _CTX_COLL))._wait()

    barrier = Barrier

    def Abort(self, errorcode=1):
        self._eng.finish_all()
        os._exit(errorcode)


def _make_world():
    rank = int(os.environ.get("PDNN_MPI_RANK", "0"))
    size = int(os.environ.get("PDNN_MPI_SIZE", "1"))
    rdv = os.environ.get("PDNN_MPI_DIR", "")
    if size > 1 and not rdv:
        raise RuntimeError("mpi4py shim: PDNN_MPI_DIR must name the rendezvous directory when PDNN_MPI_SIZE > 1")
    eng = _Engine(rank, size, rdv)
    atexit.register(eng.close)
    return Intracomm(eng)


COMM_WORLD = _make_world()
COMM_SELF = COMM_WORLD if COMM_WORLD.Get_size() == 1 else Intracomm(_Engine(0, 1, ""))


def Wtime():
    return time.time()


def Get_processor_name():
    return socket.gethostname()


def Is_initialized():
    return True


def Finalize():
    COMM_WORLD._eng.close()
