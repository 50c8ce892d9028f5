"""A few lines of protobuf wire format, enough to build api.v1.beta1 requests and read replies in the tests without
the reference's generated stubs (which cannot travel to the GPU box).  Field numbers follow
pkg/apis/manager/v1beta1/api.proto of the reference; tools/gen_wire_golden.py checks this encoder byte-for-byte
against the reference's own api_pb2 and stores golden requests under tests/golden/."""


def _varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _ld(field, payload):
    if isinstance(payload, str):
        payload = payload.encode()
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def _vi(field, v):
    return _varint(field << 3) + _varint(v)


def _s(field, text):
    return _ld(field, text) if text else b""


PARAM_TYPE = {"DOUBLE": 1, "INT": 2, "DISCRETE": 3, "CATEGORICAL": 4}
CONDITION = {"CREATED": 0, "RUNNING": 1, "SUCCEEDED": 2, "KILLED": 3, "FAILED": 4, "METRICSUNAVAILABLE": 5,
             "EARLYSTOPPED": 6, "UNKNOWN": 7}
OBJECTIVE = {"UNKNOWN": 0, "MINIMIZE": 1, "MAXIMIZE": 2}


def parameter(name, ptype, max=None, min=None, list=None, step=None):
    fs = _s(1, max or "") + _s(2, min or "") + b"".join(_ld(3, v) for v in (list or [])) + _s(4, step or "")
    body = _s(1, name) + (_vi(2, PARAM_TYPE[ptype]) if PARAM_TYPE[ptype] else b"")
    return body + _ld(3, fs)


def experiment(name, algorithm, settings, objective_type, metric, params):
    algo = _s(1, algorithm) + b"".join(_ld(2, _s(1, k) + _s(2, v)) for k, v in settings)
    t = OBJECTIVE[objective_type]
    obj = (_vi(1, t) if t else b"") + _s(3, metric)
    spec = _ld(1, b"".join(_ld(1, p) for p in params)) + _ld(2, obj) + _ld(3, algo)
    return _s(1, name) + _ld(2, spec)


def trial(name, assignments, condition="SUCCEEDED", metrics=(), start_time="", completion_time=""):
    pa = b"".join(_ld(1, _s(1, k) + _s(2, v)) for k, v in assignments)
    spec = _ld(3, pa)
    c = CONDITION[condition]
    obs = b"".join(_ld(1, _s(1, k) + _s(2, v)) for k, v in metrics)
    status = _s(1, start_time) + _s(2, completion_time) + (_vi(3, c) if c else b"") + (_ld(4, obs) if metrics else b"")
    return _s(1, name) + _ld(2, spec) + _ld(3, status)


def get_suggestions_request(exp, trials=(), current_request_number=0, total_request_number=0):
    return (_ld(1, exp) + b"".join(_ld(2, t) for t in trials)
            + (_vi(4, current_request_number) if current_request_number else b"")
            + (_vi(5, total_request_number) if total_request_number else b""))


def validate_request(exp):
    return _ld(1, exp)


# ---- decoding of GetSuggestionsReply -------------------------------------------------------------------------------
def _fields(buf):
    i, n = 0, len(buf)
    while i < n:
        key, shift = 0, 0
        while True:
            b = buf[i]; i += 1
            key |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, shift = 0, 0
            while True:
                b = buf[i]; i += 1
                v |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            yield field, v
        elif wt == 2:
            ln, shift = 0, 0
            while True:
                b = buf[i]; i += 1
                ln |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            yield field, bytes(buf[i:i + ln])
            i += ln
        else:
            raise ValueError("unexpected wire type %d" % wt)


def parse_reply(buf):
    """-> list (one per suggestion) of [(name, value), ...]"""
    out = []
    for f, payload in _fields(buf):
        if f != 1:
            continue
        assignments = []
        for g, kv in _fields(payload):
            if g != 1:
                continue
            d = dict(_fields(kv))
            assignments.append((d.get(1, b"").decode(), d.get(2, b"").decode()))
        out.append(assignments)
    return out
