"""Python mirror of the reference's public API (src/lib.rs, src/client_key.rs, src/server_key.rs,
src/ciphertext/*.rs): generate_keys, ClientKey, ServerKey, FheString / FheUsize / FheOption /
FheStringSliceVector with the same method names and argument meaning.  Server methods compile the
operation to a level program (libfhs_string) and run it on the GPU (libfhs_cuda)."""
import ctypes

import numpy as np

from .engine import Engine, FhsError, Params
from .program import Program, load_host_library


def _client_lib():
    L = load_host_library()
    if not getattr(L, "_client_bound", False):
        vp, sz = ctypes.c_void_p, ctypes.c_size_t
        L.fhs_generate_keys.argtypes = [vp, vp, sz, ctypes.c_uint64, ctypes.POINTER(vp)]
        L.fhs_client_key_destroy.argtypes = [vp]
        L.fhs_client_key_destroy.restype = None
        L.fhs_client_last_error.restype = ctypes.c_char_p
        L.fhs_client_key_info.argtypes = [vp, vp, vp, vp, vp]
        L.fhs_client_ksk.argtypes = [vp, vp]
        L.fhs_client_ksk.restype = ctypes.POINTER(ctypes.c_uint64)
        L.fhs_client_bsk.argtypes = [vp, vp]
        L.fhs_client_bsk.restype = ctypes.POINTER(ctypes.c_uint64)
        L.fhs_encrypt_blocks.argtypes = [vp, vp, sz, vp]
        L.fhs_decrypt_blocks.argtypes = [vp, vp, sz, vp]
        L.fhs_string_layout.argtypes = [vp, ctypes.c_char_p, ctypes.c_int64, vp, sz, vp]
        L._client_bound = True
    return L


class ClientKey:
    """src/client_key.rs:16-62."""

    def __init__(self, l=255, seed=0, params=None):
        self.L = _client_lib()
        h = ctypes.c_void_p()
        rc = self.L.fhs_generate_keys(ctypes.byref(params) if params is not None else None, None, int(l),
                                      int(seed), ctypes.byref(h))
        if rc != 0:
            raise FhsError(self.L.fhs_client_last_error().decode())
        self.h = h
        self.p = Params()
        nbc, nbu, ml = ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_size_t()
        self.L.fhs_client_key_info(self.h, ctypes.byref(self.p), ctypes.byref(nbc), ctypes.byref(nbu), ctypes.byref(ml))
        self.num_blocks_char, self.num_blocks_usize, self.max_len = nbc.value, nbu.value, ml.value
        self.big = self.p.glwe_dim * self.p.poly_size + 1

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.fhs_client_key_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def msg_mod(self):
        return self.p.message_modulus

    def trivial_blocks(self, values):
        delta = (1 << 63) // (self.p.message_modulus * self.p.carry_modulus)
        out = np.zeros((len(values), self.big), dtype=np.uint64)
        out[:, -1] = np.asarray([int(v) * delta for v in values], dtype=np.uint64)
        return out

    def ksk(self):
        n = ctypes.c_size_t()
        ptr = self.L.fhs_client_ksk(self.h, ctypes.byref(n))
        return np.ctypeslib.as_array(ptr, shape=(n.value,))

    def bsk(self):
        n = ctypes.c_size_t()
        ptr = self.L.fhs_client_bsk(self.h, ctypes.byref(n))
        return np.ctypeslib.as_array(ptr, shape=(n.value,))

    def encrypt_blocks(self, msgs):
        msgs = np.ascontiguousarray(msgs, dtype=np.uint64).reshape(-1)
        out = np.empty((len(msgs), self.big), dtype=np.uint64)
        if self.L.fhs_encrypt_blocks(self.h, msgs.ctypes.data, len(msgs), out.ctypes.data) != 0:
            raise FhsError("encrypt failed")
        return out

    def decrypt_blocks(self, cts):
        cts = np.ascontiguousarray(cts, dtype=np.uint64).reshape(-1, self.big)
        out = np.empty(len(cts), dtype=np.uint64)
        if self.L.fhs_decrypt_blocks(self.h, cts.ctypes.data, len(cts), out.ctypes.data) != 0:
            raise FhsError("decrypt failed")
        return out

    def _radix(self, blocks, nb):
        v = self.decrypt_blocks(blocks)
        return [int(sum(int(v[i + k]) << (2 * k) for k in range(nb))) for i in range(0, len(v), nb)]

    # client_key.rs:23-43
    def encrypt_usize(self, v):
        nb = self.num_blocks_usize
        return FheUsize(self.encrypt_blocks([(int(v) >> (2 * k)) & 3 for k in range(nb)]))

    def decrypt_usize(self, ct):
        return self._radix(ct.blocks, self.num_blocks_usize)[0] % (4 ** self.num_blocks_usize)

    def decrypt_bool(self, ct):
        return int(self.decrypt_blocks(ct.blocks)[0]) != 0

    def encrypt_str(self, s, l=None):
        """FheString::new (mod.rs:58-92); raises ValueError with the reference's messages."""
        if any(ord(c) >= 128 for c in s):
            raise ValueError("string is not ascii")
        if "\0" in s:
            raise ValueError("string contains terminator character")
        n = ctypes.c_size_t()
        pad = -1 if l is None else int(l)
        rc = self.L.fhs_string_layout(self.h, s.encode("ascii"), pad, None, 0, ctypes.byref(n))
        if rc != 0:
            raise ValueError(self.L.fhs_client_last_error().decode())
        vals = np.zeros(n.value * self.num_blocks_char, dtype=np.uint64)
        self.L.fhs_string_layout(self.h, s.encode("ascii"), pad, vals.ctypes.data, len(vals), ctypes.byref(n))
        return FheString(self.encrypt_blocks(vals))

    def decrypt_char(self, c):
        return self._radix(c.blocks, self.num_blocks_char)[0] % 256

    def decrypt_str(self, fs):
        chars = self._radix(fs.blocks, self.num_blocks_char)
        return "".join(chr(c % 256) for c in chars if c % 256 != 0)


class _DevBuf:
    """A device-resident array of result blocks (fhs_dev_alloc); freed when the last value that slices
    it goes away.  The first host access downloads the whole array once."""

    def __init__(self, engine, n_blocks):
        self.engine, self.n = engine, int(n_blocks)
        self.ptr = engine.dev_alloc(self.n)
        self._host = None

    def host(self):
        if self._host is None:
            self._host = self.engine.dev_read(self.ptr, self.n)
        return self._host

    def __del__(self):
        try:
            self.engine.dev_free(self.ptr)
        except Exception:
            pass


class _Ct:
    """Blocks of one encrypted value: a host array (fresh encryptions), or a slice of a device-resident
    result (`dev` = (_DevBuf, first block, n blocks)) that is downloaded only when `.blocks` is read --
    the result of one server operation is the operand of the next without crossing PCIe."""

    def __init__(self, blocks=None, dev=None):
        self._host = None if blocks is None else np.ascontiguousarray(blocks, dtype=np.uint64)
        self._dev = dev
        self.n_blocks = len(self._host) if self._host is not None else int(dev[2])

    @property
    def blocks(self):
        if self._host is None:
            buf, off, n = self._dev
            self._host = buf.host()[off:off + n]
        return self._host

    def is_resident(self):
        return self._dev is not None

    def _operand(self, engine, upload=False):
        """("dev", pointer, n) when the blocks are resident on this engine's GPU, else ("host", array).
        upload=True moves a host value to the device at its first use and keeps it there."""
        if self._dev is not None and not (self._dev[0].engine is engine and engine.h):
            _ = self.blocks                      # resident elsewhere: go through the host copy
            self._dev = None
        if self._dev is None and upload:
            buf = _DevBuf(engine, self.n_blocks)
            engine.dev_write(buf.ptr, self._host)
            buf._host = self._host
            self._dev = (buf, 0, self.n_blocks)
        if self._dev is not None:
            buf, off, n = self._dev
            return ("dev", buf.ptr + off * engine.big * 8, n)
        return ("host", self.blocks)


class _Out:
    """Result blocks of one program run, host or device; take(a, b) -> constructor arguments of a value."""

    def __init__(self, host=None, dev=None):
        self.host, self.dev = host, dev

    def take(self, a, b):
        if self.dev is not None:
            return dict(dev=(self.dev, a, b - a))
        return dict(blocks=self.host[a:b])

    def __len__(self):
        return self.dev.n if self.dev is not None else len(self.host)


class ServerKey:
    """src/server_key.rs:12-33: key material on the GPU + program cache.

    resident=True (default): results stay in HBM as device-resident arrays and are downloaded when the
    client first reads them (decrypt); chained operations never cross PCIe.  resident=False returns host
    arrays like the reference.  With an engine that has a communicator (Engine.comm_init*, one process
    per GPU) every operation runs sharded: all ranks call it with the same operands."""

    def __init__(self, ck, device=0, engine=None, resident=True):
        if engine is None:
            engine = Engine(device)
            engine.upload_keys(ck.p, ck.ksk(), ck.bsk(), ksk_level_order=0)
        self.engine = engine
        self.num_blocks_char, self.num_blocks_usize = ck.num_blocks_char, ck.num_blocks_usize
        self.max_len = ck.max_len
        self.resident = resident
        self.shard_min = 74
        self._progs = {}
        self._arena = None
        self.last_stats = None

    def msg_mod(self):
        return self.engine.p.message_modulus

    def trivial_blocks(self, values):
        """create_trivial_*: mask 0, body = value * Delta (mod.rs:109,266,427 -- not hidden, no noise)."""
        p = self.engine.p
        delta = (1 << 63) // (p.message_modulus * p.carry_modulus)
        out = np.zeros((len(values), self.engine.big), dtype=np.uint64)
        out[:, -1] = np.asarray([int(v) * delta for v in values], dtype=np.uint64)
        return out

    def _stage(self, parts):
        """Host operand blocks of one call, copied into a page-locked arena that lives with the engine
        (fhs_host_alloc): the H2D transfer is then a direct DMA instead of a staged copy."""
        pinned = [x.flags["C_CONTIGUOUS"] and self.engine.is_pinned(x) for x in parts]   # e.g. earlier results
        n = sum(len(x) for x, pin in zip(parts, pinned) if not pin)
        if n and (self._arena is None or self._arena.shape[0] < n):
            self._arena = self.engine.pinned_blocks(max(2 * n, 1024))
        out, o = [], 0
        for x, pin in zip(parts, pinned):
            if pin:
                out.append(x)
                continue
            self._arena[o:o + len(x)] = x
            out.append(self._arena[o:o + len(x)])
            o += len(x)
        return out

    def program(self, op, n_chars=(), n_usize=0, scalars=()):
        key = (op, tuple(n_chars), int(n_usize), tuple(scalars))
        p = self._progs.get(key)
        if p is None:
            p = Program(op, list(n_chars), n_usize, scalars, self.num_blocks_char, self.num_blocks_usize)
            self._progs[key] = p
        return p

    def execute(self, p, values):
        """Run program p on encrypted values (anything with `_operand`), in order."""
        eng = self.engine
        raw = [v._operand(eng, upload=self.resident) for v in values]
        staged = iter(self._stage([o[1] for o in raw if o[0] == "host"]))
        ops = []
        for o in raw:
            if o[0] == "host":
                ops.append(("host", next(staged)))
            elif ops and ops[-1][0] == "dev" and ops[-1][1] + ops[-1][2] * eng.big * 8 == o[1]:
                ops[-1] = ("dev", ops[-1][1], ops[-1][2] + o[2])     # adjacent slices of one result: one copy
            else:
                ops.append(o)
        rank, world = eng.comm if eng.comm else (0, 1)
        if self.resident:
            buf = _DevBuf(eng, p.n_outputs)
            self.last_stats = p.run_ex(eng, ops, out_dev=buf.ptr, rank=rank, world=world, shard_min=self.shard_min)
            out = _Out(dev=buf)
        else:
            host = eng.pinned_array(p.n_outputs)      # page-locked: direct DMA, no staged copy
            self.last_stats = p.run_ex(eng, ops, out_host=host, rank=rank, world=world, shard_min=self.shard_min)
            out = _Out(host=host)
        return _wrap(p.desc, out, self.num_blocks_usize)

    def run(self, op, strings=(), usizes=(), scalars=()):
        p = self.program(op, [s.n_chars for s in strings], len(usizes), scalars)
        return self.execute(p, list(strings) + list(usizes))


def generate_keys(l=255, seed=0, device=0):
    """src/lib.rs:44-77: (ClientKey, ServerKey) for PARAM_MESSAGE_2_CARRY_2_KS_PBS, max length l."""
    ck = ClientKey(l=l, seed=seed)
    return ck, ServerKey(ck, device)


class BooleanBlock(_Ct):
    def decrypt(self, k):
        return k.decrypt_bool(self)


class FheUsize(_Ct):
    """mod.rs:402-442."""

    def decrypt(self, k):
        return k.decrypt_usize(self)

    @staticmethod
    def new(k, v):
        """mod.rs:406: encrypt v under the client key k, num_blocks_usize blocks."""
        return k.encrypt_usize(v)

    @staticmethod
    def new_trivial(k, v):
        """mod.rs:426: trivial (unhidden) encryption made by the SERVER key."""
        nb = k.num_blocks_usize
        return FheUsize(k.trivial_blocks([(int(v) >> (2 * i)) & 3 for i in range(nb)]))

    @staticmethod
    def new_from_bool(k, b):
        """mod.rs:431: b ? 1 : 0 as a usize -- the boolean block followed by trivial zeros, no PBS."""
        return FheUsize(np.concatenate([b.blocks, k.trivial_blocks([0] * (k.num_blocks_usize - 1))]))

    @staticmethod
    def max(k):
        """mod.rs:435: msg_mod ^ num_blocks_usize - 1 (k: ClientKey or ServerKey)."""
        return k.msg_mod() ** k.num_blocks_usize - 1


class FheAsciiChar(_Ct):
    """mod.rs:31: one char = num_blocks_char radix blocks."""

    def decrypt(self, k):
        return k.decrypt_char(self)

    def _as_str(self):
        s = FheString(dev=self._dev) if self._host is None else FheString(self._host)
        return s

    def is_uppercase(self, k):      # convert.rs:14
        return k.run("char_is_uppercase", [self._as_str()])

    def is_lowercase(self, k):      # convert.rs:22
        return k.run("char_is_lowercase", [self._as_str()])

    def is_whitespace(self, k):     # trim.rs:15
        return k.run("char_is_whitespace", [self._as_str()])


class FheOption:
    """mod.rs:358-399: is_some + val; decrypt gives None when is_some is false."""

    def __init__(self, is_some, val):
        self.is_some, self.val = is_some, val

    def decrypt(self, k):
        if not self.is_some.decrypt(k):
            return None
        if isinstance(self.val, tuple):
            return tuple(v.decrypt(k) for v in self.val)
        return self.val.decrypt(k)


class FheStringSliceVector:
    """split.rs:21-231: the string + candidate (is_some, start, end) entries."""

    def __init__(self, s, entries):
        self.s, self.v = s, entries

    def _values(self):
        vals = [self.s]
        for is_some, start, end in self.v:
            vals += [is_some, start, end]
        return vals

    def _run(self, k, op, extra=()):
        p = k.program(op, [self.s.n_chars], 0, (len(self.v),))
        return k.execute(p, self._values() + list(extra))

    def len(self, k):
        """split.rs:32: number of substrings (encrypted)."""
        return self._run(k, "slices_len")

    def truncate(self, k, i):
        """split.rs:86: drop every substring from index i on (in place, like the reference)."""
        flags = self._run(k, "slices_truncate", [i])
        self.v = [(f, start, end) for f, (_, start, end) in zip(flags, self.v)]

    def reverse(self):
        """split.rs:229: reverse the order of the elements (no server work)."""
        self.v.reverse()

    def get(self, k, i):
        """split.rs:46: the substring at encrypted index i, if existent."""
        return self._run(k, "slices_get", [i])

    def decrypt(self, k):
        s = self.s.decrypt(k)
        out = []
        for is_some, start, end in self.v:
            if is_some.decrypt(k):
                a, b = start.decrypt(k), end.decrypt(k)
                out.append(s[a:b] if a <= b else "")
        return out


def _wrap(desc, out, nbu):
    t = out.take
    if desc == "bool":
        return BooleanBlock(**t(0, 1))
    if desc == "usize":
        return FheUsize(**t(0, nbu))
    if desc == "char":
        return FheAsciiChar(**t(0, len(out)))
    if desc.startswith("bools:"):
        return [BooleanBlock(**t(i, i + 1)) for i in range(int(desc.split(":")[1]))]
    if desc.startswith("str:"):
        return FheString(**t(0, len(out)))
    if desc == "opt_usize":
        return FheOption(BooleanBlock(**t(0, 1)), FheUsize(**t(1, 1 + nbu)))
    if desc.startswith("opt_str:"):
        return FheOption(BooleanBlock(**t(0, 1)), FheString(**t(1, len(out))))
    if desc.startswith("opt_pair:"):
        _, n1, _n2 = desc.split(":")
        n1 = int(n1) * 4
        return FheOption(BooleanBlock(**t(0, 1)), (FheString(**t(1, 1 + n1)), FheString(**t(1 + n1, len(out)))))
    if desc.startswith("slices:"):
        _, n_entries, _, n_chars = desc.split(":")
        p = 4 * int(n_chars)
        s = FheString(**t(0, p))
        ents = []
        for _ in range(int(n_entries)):
            ents.append((BooleanBlock(**t(p, p + 1)), FheUsize(**t(p + 1, p + 1 + nbu)),
                         FheUsize(**t(p + 1 + nbu, p + 1 + 2 * nbu))))
            p += 1 + 2 * nbu
        return FheStringSliceVector(s, ents)
    raise ValueError(desc)


class FheString(_Ct):
    """mod.rs:43-47: Vec<FheAsciiChar>, 0-terminated; blocks = [n_chars * 4][2049] u64."""

    def __init__(self, blocks=None, dev=None):
        super().__init__(blocks, dev)
        self.n_chars = self.n_blocks // 4

    def decrypt(self, k):
        return k.decrypt_str(self)

    def max_len(self):
        return self.n_chars - 1

    @staticmethod
    def new(k, s, l=None):
        """mod.rs:58: FheString::new(client key, s, pad length)."""
        return k.encrypt_str(s, l)

    @staticmethod
    def new_trivial(k, s):
        """mod.rs:96: trivial encryption by the SERVER key (same three checks, no padding)."""
        if any(ord(c) >= 128 for c in s):
            raise ValueError("string is not ascii")
        if "\0" in s:
            raise ValueError("string contains terminator character")
        if len(s) > FheString.max_len_with_key(k):
            raise ValueError("string length exceeds maximum length")
        vals = [(ord(c) >> (2 * b)) & 3 for c in s + "\0" for b in range(k.num_blocks_char)]
        return FheString(k.trivial_blocks(vals))

    @staticmethod
    def max_len_with_key(k):
        """mod.rs:240."""
        return FheUsize.max(k)

    def char_at(self, k, i):
        """mod.rs:217: the char at encrypted index i, 0 when out of bounds."""
        return k.run("char_at", [self], [i])

    def substr_end(self, k, start, end):
        """mod.rs:198: self[start..end]."""
        return k.run("substr_end", [self], [start, end])

    def substr_eq(self, k, i, s):
        """compare.rs:119: self[i..i+len(s)] == s for a CLEAR offset i."""
        return k.run("substr_eq", [self, s], (), (int(i),))


def _unary(name):
    def f(self, k):
        return k.run(name, [self])
    f.__name__ = name
    return f


def _binary(name):
    def f(self, k, s):
        return k.run(name, [self, s])
    f.__name__ = name
    return f


for _n in ("is_empty", "len", "to_uppercase", "to_lowercase", "trim", "trim_start", "trim_end",
           "split_ascii_whitespace"):
    setattr(FheString, _n, _unary(_n))
for _n in ("eq", "ne", "lt", "le", "gt", "ge", "eq_ignore_ascii_case", "contains", "find", "rfind",
           "starts_with", "ends_with", "strip_prefix", "strip_suffix", "split", "rsplit", "split_inclusive",
           "split_terminator", "rsplit_terminator", "split_once", "rsplit_once"):
    setattr(FheString, _n, _binary(_n))
FheString.substr_to = lambda self, k, index: k.run("substr_to", [self], [index])
FheString.substr_from = lambda self, k, index: k.run("substr_from", [self], [index])
FheString.splitn = lambda self, k, n, p: k.run("splitn", [self, p], [n])
FheString.rsplitn = lambda self, k, n, p: k.run("rsplitn", [self, p], [n])
# insert.rs:14-124
FheString.add = lambda self, k, s: k.run("add", [self, s])
FheString.insert = lambda self, k, index, s: k.run("insert", [self, s], [index])
FheString.repeat = lambda self, k, n, l: k.run("repeat", [self], [n], (int(l),))
FheString.replace = lambda self, k, p, s, l: k.run("replace", [self, p, s], (), (int(l),))
FheString.replacen = lambda self, k, p, s, n, l: k.run("replacen", [self, p, s], [n], (int(l),))
