"""bincode reader / writer for the reference's ciphertext types -- **RECALLED LAYOUT, UNVERIFIED**.

The reference derives serde for FheString / FheUsize (src/ciphertext/mod.rs:30,42,401) over tfhe-rs 0.5
`shortint::Ciphertext`; a user ships them with bincode 1.x defaults (little endian, fixed-width integers,
u64 lengths, u32 enum variants).  INTEGRATION.md writes the field order down; this module is that table as
code, so that ciphertexts serialized by the reference can be fed to the engine (`blocks`) and results
written back, and so that the first box with a Rust toolchain can falsify the recollection with ONE file:
`parse_fhe_string(open("s.bin","rb").read())` either consumes the file exactly or raises.

Per shortint block:  u64 len (= lwe_size) | len x u64 (mask || body) | u128 modulus (0 = native 2^64) |
u64 scalar_bits | u64 degree | u64 noise_level | u64 message_modulus | u64 carry_modulus | u32 pbs_order.
"""
import struct

import numpy as np

BLOCK_META = struct.Struct("<QQQQQQQI")     # modulus lo, modulus hi, scalar_bits, degree, noise_level, msg_mod, carry_mod, pbs_order


class WireError(ValueError):
    pass


def _u64(buf, off):
    if off + 8 > len(buf):
        raise WireError("truncated input")
    return struct.unpack_from("<Q", buf, off)[0], off + 8


def parse_block(buf, off=0):
    """-> (u64[lwe_size] array, meta dict, new offset)"""
    n, off = _u64(buf, off)
    if n == 0 or n > (1 << 20) or off + 8 * n + BLOCK_META.size > len(buf):
        raise WireError(f"implausible LWE size {n}")
    ct = np.frombuffer(buf, dtype="<u8", count=n, offset=off).copy()
    off += 8 * n
    mlo, mhi, bits, degree, noise, msg, carry, order = BLOCK_META.unpack_from(buf, off)
    off += BLOCK_META.size
    if (mlo, mhi) != (0, 0) or bits != 64:
        raise WireError(f"not a native 2^64 ciphertext modulus: ({mlo}, {mhi}), {bits} bits")
    if order not in (0, 1):
        raise WireError(f"bad PBSOrder variant {order}")
    return ct, dict(degree=degree, noise_level=noise, message_modulus=msg, carry_modulus=carry, pbs_order=order), off


def parse_radix(buf, off=0):
    """RadixCiphertext { blocks: Vec<Ciphertext> } -> (u64[n_blocks][lwe_size], [meta], new offset)"""
    n, off = _u64(buf, off)
    if n > (1 << 24):
        raise WireError(f"implausible block count {n}")
    blocks, metas = [], []
    for _ in range(n):
        ct, meta, off = parse_block(buf, off)
        blocks.append(ct)
        metas.append(meta)
    return (np.stack(blocks) if blocks else np.zeros((0, 0), np.uint64)), metas, off


def parse_fhe_string(buf):
    """FheString(Vec<FheAsciiChar(RadixCiphertext)>) -> u64[n_chars * blocks_per_char][lwe_size]; the whole
    buffer must be consumed."""
    n_chars, off = _u64(buf, 0)
    if n_chars == 0 or n_chars > (1 << 20):
        raise WireError(f"implausible char count {n_chars}")
    chars = []
    for _ in range(n_chars):
        b, metas, off = parse_radix(buf, off)
        for m in metas:
            if m["pbs_order"] != 0 or m["message_modulus"] != 4 or m["carry_modulus"] != 4:
                raise WireError(f"block is not a PARAM_MESSAGE_2_CARRY_2_KS_PBS ciphertext: {m}")
        chars.append(b)
    if off != len(buf):
        raise WireError(f"{len(buf) - off} trailing bytes: the recalled layout does not match this file")
    return np.concatenate(chars)


def parse_fhe_usize(buf):
    b, _, off = parse_radix(buf, 0)
    if off != len(buf):
        raise WireError(f"{len(buf) - off} trailing bytes")
    return b


def dump_block(ct, degree=3, noise_level=1, message_modulus=4, carry_modulus=4, pbs_order=0):
    ct = np.ascontiguousarray(ct, dtype="<u8")
    return (struct.pack("<Q", ct.size) + ct.tobytes() +
            BLOCK_META.pack(0, 0, 64, degree, noise_level, message_modulus, carry_modulus, pbs_order))


def dump_radix(blocks, degree=3):
    return struct.pack("<Q", len(blocks)) + b"".join(dump_block(b, degree=degree) for b in blocks)


def dump_fhe_string(blocks, blocks_per_char=4):
    """u64[n_chars * blocks_per_char][lwe_size] -> bincode bytes of an FheString."""
    blocks = np.asarray(blocks, dtype=np.uint64)
    n_chars = len(blocks) // blocks_per_char
    return struct.pack("<Q", n_chars) + b"".join(
        dump_radix(blocks[i * blocks_per_char:(i + 1) * blocks_per_char]) for i in range(n_chars))
