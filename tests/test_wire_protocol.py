"""Wire layer of the node: the hand-written args.proto codec and SP REQ client
against independent implementations — Google's protobuf runtime and a REP server on raw
sockets (tests/sp_server.py).  No GPU: gsmfretda_b200/host/wire_probe links no CUDA library."""
import os
import struct
import subprocess
import time

import numpy as np
import pytest

from sp_server import PB, ParamServer, RepServer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE = os.path.join(ROOT, "gsmfretda_b200", "host", "wire_probe")


@pytest.fixture(scope="module", autouse=True)
def _built():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "gsmfretda_b200", "host"), PROBE], check=True)


def probe(*args, env=None, timeout=60):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([PROBE, *map(str, args)], capture_output=True, text=True, timeout=timeout, env=e)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    return r.stdout.strip().split("\n")


def fhex(x):
    return "0x%08x" % struct.unpack("<I", struct.pack("<f", x))[0]


def decode_fields(lines):
    return dict(l.split("=", 1) for l in lines)


# ---- codec: same bytes as libprotobuf for the same fields --------------------------------------
CASES = [
    ("p_cap", dict(cap=7733)),
    ("p_cap", dict(cap=0)),                       # proto3: default value -> empty message
    ("p_n", dict(s_n=3)),
    ("p_sid", dict(sid=-1)),                      # 10-byte varint
    ("p_sid", dict(sid=2147483647)),
    ("p_sid", dict(sid=-2147483648)),
    ("p_ga", dict(idx=5, start=0, stop=-1, bestfv=3.1e32, params=[0.25, 0.75, 500.0, 500.0, 3.0, 3.0])),
    ("p_ga", dict(idx=-2, bestfv=0.0)),
    ("p_ga", dict(idx=0, start=10, stop=20000, bestfv=1.25, params=[])),
    ("p_str", dict(histNum=200, str="00ff-12-140-0-3", ohist=[float(i % 17) for i in range(200)])),
    ("p_str", dict(histNum=0, str="node")),
    ("res", dict(ridx=41, e=1.7320508, hist=True, idx="abcd-1-2-0-0", shist=[i / 4.0 for i in range(200)])),
    ("res", dict(ridx=0, e=float("inf"), hist=False, idx="x")),
    ("res", dict(ridx=3, e=float("nan"), hist=False, idx="")),
    # batch extension: chi-square of every vector + offset of the best one
    ("res", dict(ridx=8, e=0.5, hist=True, idx="n", shist=[1.0, 2.0], es=[2.5, 0.5, 7.0], best=1)),
    ("res", dict(ridx=8, e=2.5, hist=False, idx="n", es=[2.5, 9.5], best=0)),
]


@pytest.mark.parametrize("name,fields", CASES)
def test_encode_matches_protobuf_runtime(name, fields):
    want = PB[name](**fields).SerializeToString()
    args = []
    for k, v in fields.items():
        if isinstance(v, list):
            if v:
                args.append("%s=%s" % (k, ",".join(fhex(x) for x in v)))
        elif isinstance(v, bool):
            args.append("%s=%d" % (k, int(v)))
        elif isinstance(v, float):
            args.append("%s=%s" % (k, fhex(v)))
        else:
            args.append("%s=%s" % (k, v))
    got = bytes.fromhex(probe("encode", name, *args)[0])
    assert got == want


@pytest.mark.parametrize("name,fields", CASES)
def test_decode_what_protobuf_runtime_writes(name, fields):
    m = PB[name](**fields)
    raw = m.SerializeToString()
    d = decode_fields(probe("decode", name, raw.hex() or "00"[:0] or ""))
    assert d["ok"] == "1"
    for k, v in fields.items():
        if isinstance(v, list):
            got = [x for x in d[k].split(",") if x]
            assert got == [fhex(x) for x in v]
        elif isinstance(v, bool):
            assert d[k] == str(int(v))
        elif isinstance(v, float):
            assert d[k] == fhex(v)
        else:
            assert d[k] == str(v)


def test_decode_unpacked_repeated_unknown_fields_and_last_wins():
    # params as five separate fixed32 fields (an old encoder), an unknown varint field 9, an
    # unknown length-delimited field 10, idx written twice: protobuf semantics
    raw = b""
    raw += bytes([0x08, 0x01])
    for x in (1.0, 2.0, 3.0):
        raw += bytes([0x2D]) + struct.pack("<f", x)
    raw += bytes([0x48, 0x96, 0x01])
    raw += bytes([0x52, 0x03]) + b"abc"
    raw += bytes([0x2A, 0x08]) + struct.pack("<ff", 4.0, 5.0)
    raw += bytes([0x08, 0x07])
    ref = PB["p_ga"]()
    ref.ParseFromString(raw)
    d = decode_fields(probe("decode", "p_ga", raw.hex()))
    assert d["ok"] == "1" and d["idx"] == "7" == str(ref.idx)
    assert d["params"].split(",") == [fhex(x) for x in ref.params] == [fhex(x) for x in (1, 2, 3, 4, 5)]


@pytest.mark.parametrize("raw", [b"\x08", b"\x2a\x08\x00\x00", b"\x25\x00\x00", b"\x2a\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\x01",
                                 b"\x0b\x08\x01\x0c"])
def test_decode_rejects_truncated_and_group_fields(raw):
    d = decode_fields(probe("decode", "p_ga", raw.hex()))
    assert d["ok"] == "0"


# ---- SP framing: REQ client against the REP server on raw sockets ------------------------------
def test_req_rep_roundtrip_and_frame_bytes():
    srv = RepServer(lambda body: b"re:" + body)
    try:
        payloads = [b"c" + PB["p_cap"](cap=123).SerializeToString(), b"k", b"x" * 50000]
        out = probe("req", srv.url, *[p.hex() for p in payloads])
        assert [bytes.fromhex(o) for o in out] == [b"re:" + p for p in payloads]
        ins = [f for d, f in srv.frames if d == "in"]
        assert len(ins) == 3
        ids = [struct.unpack(">I", f[:4])[0] for f in ins]
        assert all(i & 0x80000000 for i in ids), "request id carries the end-of-backtrace bit"
        assert ids[1] == ((ids[0] + 1) | 0x80000000) and ids[2] == ((ids[1] + 1) | 0x80000000)
        assert [f[4:] for f in ins] == payloads
    finally:
        srv.close()


def test_req_rep_over_ipc(tmp_path):
    url = "ipc://" + str(tmp_path / "sp.sock")
    srv = RepServer(lambda body: body[::-1], url)
    try:
        out = probe("req", url, b"hello".hex(), b"".hex() or "")
        assert bytes.fromhex(out[0]) == b"olleh"
    finally:
        srv.close()


def test_request_is_sent_again_when_the_connection_dies():
    srv = RepServer(lambda body: b"ok" + body)
    srv.kill_next = 1          # first request: the server closes the connection without replying
    try:
        out = probe("req", srv.url, b"abc".hex(), env={"WIRE_RECONNECT_MS": "50"})
        assert bytes.fromhex(out[0]) == b"okabc"
        ins = [f for d, f in srv.frames if d == "in"]
        assert len(ins) == 2 and ins[0] == ins[1], "same id, same body on the new connection"
    finally:
        srv.close()


def test_request_is_sent_again_after_the_resend_interval():
    srv = RepServer(lambda body: b"ok" + body)
    srv.drop_next = 1          # first copy is swallowed
    try:
        t0 = time.time()
        out = probe("req", srv.url, b"abc".hex(), env={"WIRE_RESEND_MS": "300"})
        assert bytes.fromhex(out[0]) == b"okabc"
        assert time.time() - t0 >= 0.25
        ins = [f for d, f in srv.frames if d == "in"]
        assert len(ins) == 2 and ins[0] == ins[1]
    finally:
        srv.close()


def test_client_waits_for_a_server_that_starts_late():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    url = "tcp://127.0.0.1:%d" % port
    p = subprocess.Popen([PROBE, "req", url, b"late".hex()], stdout=subprocess.PIPE, text=True)
    time.sleep(0.6)                      # the client is retrying every 200 ms (streamWorker.cpp:53-54)
    srv = RepServer(lambda body: body.upper(), url)
    try:
        out, _ = p.communicate(timeout=20)
        assert bytes.fromhex(out.strip()) == b"LATE"
    finally:
        srv.close()


def test_receive_timeout_is_reported():
    srv = RepServer(lambda body: body)
    srv.drop_next = 10
    try:
        r = subprocess.run([PROBE, "req", srv.url, b"x".hex()], capture_output=True, text=True, timeout=20,
                           env=dict(os.environ, WIRE_RCVTIMEO_MS="300"))
        assert r.returncode == 1 and "recv-error" in r.stdout
    finally:
        srv.close()


# ---- the four exchanges, message level (what the node will say, said by hand) -------------------
def test_param_server_conversation():
    vecs = [[0.25, 0.75, 500, 500, 3, 3], [0.3, 0.7, 100, 200, 2, 2]]
    srv = ParamServer(2, vecs, bestfv=5.0, idle_first=1)
    try:
        c = b"c" + PB["p_cap"](cap=7733).SerializeToString()
        p1 = b"p" + PB["p_str"](str="node-0", histNum=4, ohist=[1, 2, 3, 4]).SerializeToString()
        p2 = b"p" + PB["p_str"](str="node-0").SerializeToString()
        k = b"k" + PB["p_sid"](sid=0).SerializeToString()
        r0 = b"r" + PB["res"](ridx=0, e=4.5, hist=True, idx="node-0", shist=[0.5, 1.5]).SerializeToString()
        r1 = b"r" + PB["res"](ridx=1, e=6.5, hist=False, idx="node-0").SerializeToString()
        out = [bytes.fromhex(o) for o in probe("req", srv.url, *[m.hex() for m in (c, p1, p2, k, r0, p2, r1, p2)])]
        assert PB["p_n"].FromString(out[0]).s_n == 2
        assert PB["p_ga"].FromString(out[1]).idx == -1                       # nothing yet: ask again
        ga = PB["p_ga"].FromString(out[2])
        assert ga.idx == 0 and list(ga.params) == vecs[0] and ga.bestfv == 5.0 and ga.stop == -1
        assert PB["p_sid"].FromString(out[4]).sid == 0
        assert PB["p_ga"].FromString(out[5]).idx == 1
        assert PB["p_sid"].FromString(out[6]).sid == -1                      # list done: leave
        assert PB["p_ga"].FromString(out[7]).idx == -2
        assert srv.caps == [7733] and srv.nodes["node-0"]["ohist"] == [1, 2, 3, 4]
        assert srv.results[0]["shist"] == [0.5, 1.5] and srv.results[1]["hist"] is False
        assert np.float32(srv.results[0]["e"]) == np.float32(4.5)
    finally:
        srv.close()
