"""Second, independent restatement of the routing SPEC (DESIGN.md "SPEC") in pure Python.

TEST INFRASTRUCTURE ONLY.  Written from the prose spec, byte at a time with Python
integers, to cross-check oracle/pk_oracle.c on small inputs (it is ~10^5 times slower).
Tool classification follows internal/service/mock.go:32-64; everything else is
build-defined (parity unpinned by the reference).
"""
M32 = 0xFFFFFFFF
M64 = 0xFFFFFFFFFFFFFFFF
P1, P2, P5 = 0x9E3779B1, 0x85EBCA77, 0x165667B1
NAME_MUL = 0xD6E8FEB86659FD93
FMIX_C1 = 0xFF51AFD7ED558CCD


def rotl(x, r):
    return ((x << r) | (x >> (32 - r))) & M32


def fmix64(x):
    x ^= x >> 33
    x = (x * 0xFF51AFD7ED558CCD) & M64
    x ^= x >> 33
    x = (x * 0xC4CEB9FE1A85EC53) & M64
    x ^= x >> 33
    return x


def tool_class(tools, default_class, name: bytes):
    """tools: [(name_bytes, class_id)].  Go string switch: exact, case-sensitive."""
    for nm, cid in tools:
        if nm == name:
            return cid, False
    return default_class, True


def is_letter(b):
    return 0x41 <= b <= 0x5A or 0x61 <= b <= 0x7A


def est_tokens(p: bytes):
    tokens = 0
    for j, b in enumerate(p):
        if b == 0x00 or b == 0x20 or 0x80 <= b <= 0xBF:
            continue                                  # NUL, space and UTF-8 continuation bytes
        if is_letter(b):
            if j > 0 and is_letter(p[j - 1]) and j % 8 != 0:
                continue                              # letter run continues
            tokens += 1
        else:
            tokens += 1                               # lead byte, digit, punctuation, control, DEL
    return tokens, any(b >= 0x80 for b in p)


def payload_sum(p: bytes):
    nblk = (len(p) + 15) // 16
    padded = p + b"\0" * (nblk * 16 - len(p))
    stripes = {}                                      # q -> state, insertion in block order
    width = 8 if nblk <= 16 else 32                   # SPEC v3: payloads of at most 16 blocks have 8 stripes
    for k in range(nblk):
        seg, r = divmod(k, 1024)
        q = (seg * 32 + (r % width)) & M32
        if q not in stripes:
            base = (q * P5) & M32
            stripes[q] = [(P1 + P2 + base) & M32, (P2 + base) & M32, base, (base - P1) & M32]
        a = stripes[q]
        for j in range(4):
            w = int.from_bytes(padded[16 * k + 4 * j:16 * k + 4 * j + 4], "little")
            a[j] = (rotl((a[j] + w * P2) & M32, 13) * P1) & M32
    total = 0
    for a in stripes.values():
        lo = (rotl(a[0], 1) + rotl(a[1], 7) + rotl(a[2], 12) + rotl(a[3], 18)) & M32
        hi = (rotl(a[0], 24) + rotl(a[1], 3) + rotl(a[2], 19) + rotl(a[3], 10)) & M32
        total = (total + fmix64((hi << 32) | lo)) & M64
    return total


def name_sum(name: bytes):
    """Name as little-endian 32-bit words (last one zero padded); word j adds mixn((j+1) << 32 | word)."""
    total = 0
    for j in range((len(name) + 3) // 4):
        w = int.from_bytes(name[4 * j:4 * j + 4].ljust(4, b"\0"), "little")
        x = ((j + 1) << 32) | w
        x = (x * NAME_MUL) & M64
        x ^= x >> 32
        x = (x * FMIX_C1) & M64
        total = (total + x) & M64
    return total


def route_hash(p: bytes, name: bytes):
    s = (payload_sum(p) + name_sum(name) + (len(name) << 32) + (len(p) & M32)) & M64
    return fmix64(s)


def select_provider(providers, h, tokens, cls):
    """providers: [(max_tokens, weight, tool_mask, reserved)] -> index or 0xFF.
    Weighted choice among the eligible providers, driven by the top 32 bits of the hash."""
    eligible = [(i, w) for i, (mt, w, mask, _) in enumerate(providers)
                if w != 0 and (mask >> cls) & 1 and tokens <= mt]
    total = sum(w for _, w in eligible)
    if total == 0:
        return 0xFF
    x = ((h >> 32) * total) >> 32
    running = 0
    for i, w in eligible:
        running += w
        if running > x:
            return i
    raise AssertionError("unreachable")


def route_one(tools, default_class, providers, name: bytes, p: bytes):
    cls, unknown = tool_class(tools, default_class, name)
    tokens, nonascii = est_tokens(p)
    h = route_hash(p, name)
    prov = select_provider(providers, h, tokens, cls)
    flags = (2 if prov == 0xFF else 0) | (4 if len(p) == 0 else 0) | (8 if unknown else 0)
    return dict(byte_len=len(p), est_tokens=tokens, route_hash=h, tool_class=cls, provider=prov,
                flags=flags, status=200)
