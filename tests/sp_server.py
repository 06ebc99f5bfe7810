"""A stand-in for pySMFRETda's pdaServ.py (absent from the reference checkout), written from the
client side of the conversation (/root/reference/src/streamWorker.cpp:83-238) and from the two
nanomsg wire documents (sp-tcp-mapping-01, sp-request-reply-01).  Test infrastructure only.

It is deliberately an INDEPENDENT implementation of both layers the node speaks:
  * REP side of nanomsg's scalability protocols over raw TCP / unix sockets (Python `socket`);
  * args.proto through Google's protobuf runtime, the message classes built from a descriptor
    that restates /root/reference/3rdparty/protobuf/args.proto:1-46 field by field.
so that a node that converses with it has byte-compatible framing and encoding.

`ParamServer` hands out a fixed list of parameter vectors (one per 'p' request), collects the
'r' results, answers keep-alives, and tells every stream to leave once the list is done — the
GA loop of the real server reduced to "evaluate these vectors".
"""
import os
import socket
import struct
import threading

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto


def _build_messages():
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name = "args.proto"
    fdp.package = "gSMFRETda.pb"
    fdp.syntax = "proto3"

    def msg(name, fields):
        m = fdp.message_type.add()
        m.name = name
        for fname, num, typ, rep in fields:
            f = m.field.add()
            f.name = fname
            f.number = num
            f.type = typ
            f.label = _F.LABEL_REPEATED if rep else _F.LABEL_OPTIONAL
            if rep:
                f.options.packed = True

    # args.proto:4-13
    msg("p_ga", [("idx", 1, _F.TYPE_INT32, 0), ("start", 2, _F.TYPE_INT32, 0), ("stop", 3, _F.TYPE_INT32, 0),
                 ("bestfv", 4, _F.TYPE_FLOAT, 0), ("params", 5, _F.TYPE_FLOAT, 1)])
    msg("p_n", [("s_n", 1, _F.TYPE_INT32, 0)])                       # :15-17
    msg("p_sid", [("sid", 1, _F.TYPE_INT32, 0)])                     # :22-27
    msg("p_cap", [("cap", 1, _F.TYPE_INT32, 0)])                     # :29-31
    msg("p_str", [("histNum", 1, _F.TYPE_INT32, 0), ("str", 2, _F.TYPE_STRING, 0),
                  ("ohist", 3, _F.TYPE_FLOAT, 1)])                   # :33-37
    # :39-45, plus the batch extension of gsmfretda_b200/host/args_pb.hpp (es = 6, best = 7)
    msg("res", [("ridx", 1, _F.TYPE_INT32, 0), ("e", 2, _F.TYPE_FLOAT, 0), ("hist", 3, _F.TYPE_BOOL, 0),
                ("idx", 4, _F.TYPE_STRING, 0), ("shist", 5, _F.TYPE_FLOAT, 1), ("es", 6, _F.TYPE_FLOAT, 1),
                ("best", 7, _F.TYPE_INT32, 0)])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    return {n: message_factory.GetMessageClass(pool.FindMessageTypeByName("gSMFRETda.pb." + n))
            for n in ("p_ga", "p_n", "p_sid", "p_cap", "p_str", "res")}


PB = _build_messages()

NN_REQ, NN_REP = 0x30, 0x31


def _recv_exact(conn, n):
    buf = b""
    while len(buf) < n:
        chunk = conn.recv(n - len(buf))
        if not chunk:
            raise ConnectionError("peer closed")
        buf += chunk
    return buf


class RepServer:
    """REP socket: accepts any number of REQ peers, hands every request body to `handler(body)`
    and sends its return value back with the request's backtrace."""

    def __init__(self, handler, url=None):
        self.handler = handler
        self.ipc = bool(url) and url.startswith("ipc://")
        if self.ipc:
            path = url[6:]
            if os.path.exists(path):
                os.unlink(path)
            self.lsock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            self.lsock.bind(path)
            self.url = url
        else:
            self.lsock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
            self.lsock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            port = int(url.rsplit(":", 1)[1]) if url else 0
            self.lsock.bind(("127.0.0.1", port))
            self.url = "tcp://127.0.0.1:%d" % self.lsock.getsockname()[1]
        self.lsock.listen(64)
        self.stop = False
        self.conns = []
        self.frames = []          # (direction, raw bytes) of everything after the handshakes
        self.drop_next = 0        # test hook: swallow this many requests without a reply
        self.kill_next = 0        # test hook: close the connection instead of replying
        self.lock = threading.Lock()
        self.thread = threading.Thread(target=self._accept, daemon=True)
        self.thread.start()

    def _accept(self):
        self.lsock.settimeout(0.2)
        while not self.stop:
            try:
                conn, _ = self.lsock.accept()
            except socket.timeout:
                continue
            except OSError:
                break
            self.conns.append(conn)
            threading.Thread(target=self._serve, args=(conn,), daemon=True).start()

    def _serve(self, conn):
        try:
            conn.sendall(bytes([0, ord("S"), ord("P"), 0, 0, NN_REP, 0, 0]))
            hello = _recv_exact(conn, 8)
            if hello[:4] != b"\x00SP\x00" or ((hello[4] << 8) | hello[5]) != NN_REQ:
                conn.close()
                return
            while not self.stop:
                if self.ipc:
                    if _recv_exact(conn, 1) != b"\x01":
                        break
                (n,) = struct.unpack(">Q", _recv_exact(conn, 8))
                payload = _recv_exact(conn, n)
                with self.lock:
                    self.frames.append(("in", payload))
                # backtrace: 32-bit words up to and including the first with the top bit set
                k = 0
                while True:
                    (w,) = struct.unpack(">I", payload[k:k + 4])
                    k += 4
                    if w & 0x80000000:
                        break
                trace, body = payload[:k], payload[k:]
                with self.lock:
                    if self.kill_next > 0:
                        self.kill_next -= 1
                        conn.close()
                        return
                    if self.drop_next > 0:
                        self.drop_next -= 1
                        continue
                reply = self.handler(body)
                out = trace + reply
                with self.lock:
                    self.frames.append(("out", out))
                conn.sendall((b"\x01" if self.ipc else b"") + struct.pack(">Q", len(out)) + out)
        except (ConnectionError, OSError, struct.error):
            pass
        finally:
            try:
                conn.close()
            except OSError:
                pass

    def close(self):
        self.stop = True
        try:
            self.lsock.close()
        except OSError:
            pass
        for c in self.conns:
            try:
                c.close()
            except OSError:
                pass
        if self.ipc and os.path.exists(self.url[6:]):
            os.unlink(self.url[6:])


class ParamServer:
    """The four exchanges of the protocol over a RepServer."""

    def __init__(self, n_states, vectors, bestfv=3.0e32, start=0, stop=-1, url=None, idle_first=0, batch=1):
        self.n_states = n_states
        self.batch = int(batch)          # vectors per 'p' reply (> 1: the batch extension)
        self.raw = []                    # every 'r' body as received (byte-level checks)
        self.vectors = [list(map(float, v)) for v in vectors]
        self.bestfv = float(bestfv)
        self.start, self.stop = start, stop
        self.next = 0
        self.idle_first = idle_first     # answer the first k 'p' requests with idx = -1
        self.caps = []
        self.nodes = {}                  # node id -> {"histNum": .., "ohist": [...]}
        self.results = {}                # ridx -> dict(e, hist, idx, shist)
        self.keepalives = []
        self.lock = threading.Lock()
        self.done = threading.Event()
        self.rep = RepServer(self._handle, url)
        self.url = self.rep.url

    def _handle(self, body):
        op, data = body[:1], body[1:]
        with self.lock:
            if op == b"c":                                    # streamWorker.cpp:83-105
                m = PB["p_cap"]()
                m.ParseFromString(data)
                self.caps.append(m.cap)
                return PB["p_n"](s_n=self.n_states).SerializeToString()
            if op == b"p":                                    # streamWorker.cpp:106-158
                m = PB["p_str"]()
                m.ParseFromString(data)
                node = self.nodes.setdefault(m.str, {"histNum": 0, "ohist": None, "requests": 0})
                node["requests"] += 1
                if m.histNum:
                    node["histNum"] = m.histNum
                    node["ohist"] = list(m.ohist)
                if self.idle_first > 0:
                    self.idle_first -= 1
                    return PB["p_ga"](idx=-1, bestfv=self.bestfv).SerializeToString()
                if self.next >= len(self.vectors):
                    return PB["p_ga"](idx=-2, bestfv=self.bestfv).SerializeToString()
                i = self.next
                take = min(self.batch, len(self.vectors) - i)
                self.next += take
                flat = [x for v in self.vectors[i:i + take] for x in v]
                return PB["p_ga"](idx=i, start=self.start, stop=self.stop, bestfv=self.bestfv,
                                  params=flat).SerializeToString()
            if op == b"k":                                    # streamWorker.cpp:165-181
                m = PB["p_sid"]()
                m.ParseFromString(data)
                self.keepalives.append(m.sid)
                return PB["p_sid"](sid=m.sid).SerializeToString()
            if op == b"r":                                    # streamWorker.cpp:207-235
                m = PB["res"]()
                m.ParseFromString(data)
                self.raw.append(bytes(data))
                if len(m.es) > 1:       # one message for a whole batch: chi-square per vector, histogram of the best
                    for k, e in enumerate(m.es):
                        self.results[m.ridx + k] = {"e": e, "hist": m.hist and k == m.best, "idx": m.idx,
                                                    "shist": list(m.shist) if k == m.best else []}
                else:
                    self.results[m.ridx] = {"e": m.e, "hist": m.hist, "idx": m.idx, "shist": list(m.shist)}
                if m.hist and m.e < self.bestfv:
                    pass   # a GA server would adopt the new best value here; fixed for the tests
                finished = len(self.results) >= len(self.vectors)
                if finished:
                    self.done.set()
                # sid == -1 makes the stream leave (streamWorker.cpp:229-230)
                return PB["p_sid"](sid=-1 if self.next >= len(self.vectors) else 0).SerializeToString()
        return b""

    def close(self):
        self.rep.close()
