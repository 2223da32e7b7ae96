"""Helpers shared by the CPU and GPU string tests: the reference's data layout in the clear."""
WS = "\t\n\x0b\x0c\r "   # Rust char::is_whitespace restricted to ASCII (trim.rs:15-25: 9..13 or 32)


def enc_chars(s, pad=None):
    """FheString::new (mod.rs:58-92): chars + terminator, optionally padded to pad+1 chars."""
    b = [ord(c) for c in s] + [0]
    if pad is not None:
        assert pad >= len(s)
        b += [0] * (pad + 1 - len(b))
    return b


def to_blocks(chars, nb=4):
    return [(c >> (2 * k)) & 3 for c in chars for k in range(nb)]


def usize_blocks(v, nb=4):
    return [(v >> (2 * k)) & 3 for k in range(nb)]


def from_blocks(blocks, nb=4):
    return [sum(blocks[i + k] << (2 * k) for k in range(nb)) for i in range(0, len(blocks), nb)]


def dec_str(blocks):
    """FheString::decrypt (mod.rs:121-129): drop every zero char."""
    return "".join(chr(c) for c in from_blocks(blocks) if c != 0)


def decode(desc, out, nbu=4):
    """Decode program outputs (clear block values) according to fhs_prog_output_desc."""
    if desc == "bool":
        return bool(out[0])
    if desc == "usize":
        return from_blocks(out, nbu)[0]
    if desc == "char":
        return from_blocks(out)[0]
    if desc.startswith("bools:"):
        return [bool(b) for b in out]
    if desc.startswith("str:"):
        return dec_str(out)
    if desc == "opt_usize":
        return from_blocks(out[1:], nbu)[0] if out[0] else None
    if desc.startswith("opt_str:"):
        return dec_str(out[1:]) if out[0] else None
    if desc.startswith("opt_pair:"):
        _, n1, n2 = desc.split(":")
        n1 = int(n1)
        return (dec_str(out[1:1 + 4 * n1]), dec_str(out[1 + 4 * n1:])) if out[0] else None
    if desc.startswith("slices:"):
        # slices:<n_entries>:str:<n_chars>  -> [string blocks][(is_some, start, end)]*
        _, n_entries, _, n_chars = desc.split(":")
        n_entries, n_chars = int(n_entries), int(n_chars)
        sb = out[: 4 * n_chars]
        chars = from_blocks(sb)
        s = "".join(chr(c) for c in chars if c != 0)   # split.rs:208-226 decrypts the string, then slices
        res = []
        p = 4 * n_chars
        for _ in range(n_entries):
            is_some = out[p]
            start = from_blocks(out[p + 1: p + 1 + nbu], nbu)[0]
            end = from_blocks(out[p + 1 + nbu: p + 1 + 2 * nbu], nbu)[0]
            p += 1 + 2 * nbu
            if is_some:
                res.append(s[start:end] if start <= end else "")
        return res
    raise ValueError(desc)
