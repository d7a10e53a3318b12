"""Generates tests/golden/route_golden.json.

    python tests/golden/make_golden.py

The reference has no tests, golden vectors or fixtures, so:
  * "tool_class" rows are written BY HAND from the literals of the Go switch at
    internal/service/mock.go:32-64 (names at :33, :37, :52, default at :60) — they are the
    only reference-pinned expectations;
  * every other expected value is computed with oracle/pk_spec_py.py, the pure-Python
    restatement of the written SPEC — so the fixture pins the C oracle (and, on the GPU,
    the kernels) to an implementation that shares no code with either.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import pk_spec_py as S  # noqa: E402

TOOLS = [(b"example_tool", 0), (b"struct_tool", 1), (b"file_tool", 2)]       # mock.go:33,37,52
DEFAULT_CLASS = 3                                                             # mock.go:60
MAXTOK = [4096, 8192, 8192, 32768, 32768, 131072, 131072, 1048576]
PROVIDERS = [(MAXTOK[p], p + 1, 0x7 if p % 4 == 3 else 0xF, 0) for p in range(8)]

# hand-written: (name, expected class, takes the default branch?)
TOOL_CLASS_ROWS = [
    (b"example_tool", 0, False), (b"struct_tool", 1, False), (b"file_tool", 2, False),
    (b"", 3, True), (b"e", 3, True), (b"Example_tool", 3, True), (b"EXAMPLE_TOOL", 3, True),
    (b"example_tool ", 3, True), (b" example_tool", 3, True), (b"example_tool\x00", 3, True),
    (b"example_too", 3, True), (b"example_tools", 3, True), (b"example-tool", 3, True),
    (b"struct_too", 3, True), (b"struct_tool\n", 3, True), (b"file_tool_", 3, True), (b"file", 3, True),
    (b"file_tool\x00\x00", 3, True), (b"unknown_tool", 3, True), (b"x" * 65, 3, True),
    (b"example_tool" * 6, 3, True), ("outil_é".encode(), 3, True), ("例".encode(), 3, True),
    (b"struct_tool_struct_tool_struct_tool_struct_tool_struct_tool_12345", 3, True),
]


def pattern(n, a=131, b=7):
    """Deterministic filler for the larger vectors (kept out of the JSON)."""
    return bytes(((i * a + i // b) % 251) for i in range(n))


def text(n):
    words = [b"route", b"the", b"Request", b"to", b"a", b"provider", b"internationalization", b"42",
             b"{\"k\":1}", "naïve".encode(), "例え".encode(), b"x_y", b"\n", b"END."]
    out = bytearray()
    i = 0
    while len(out) < n:
        out += words[(i * 5 + i // 3) % len(words)] + b" "
        i += 1
    return bytes(out[:n])


def small_vectors():
    v = [b"", b" ", b"a", b"ab", b"abcdefgh", b"abcdefghi", b"abcdefg h", b"1234", b"a1b2", b"\x00\x01\x1f\x20\x7f",
         bytes(range(256)), bytes(range(255, -1, -1)), b"The quick brown fox, jumps over 13 lazy dogs!\n",
         "café 例える \U0001F600 ok".encode(), b"\xff" * 17, b"\x80" * 33, b"A" * 15, b"A" * 16,
         b"A" * 17, b"z" * 64, b"Zz" * 100, b"a b " * 9, (b"word " * 103)[:511], (b"word " * 103)[:512],
         (b"word " * 103)[:513]]
    return v


def build():
    rows = []
    for name, cls, unknown in TOOL_CLASS_ROWS:
        got = S.tool_class(TOOLS, DEFAULT_CLASS, name)
        assert got == (cls, unknown), (name, got)           # the python restatement agrees with the hand-written rows
        rows.append({"name_hex": name.hex(), "tool_class": cls, "unknown": unknown})

    vectors = []
    names = [b"example_tool", b"struct_tool", b"file_tool", b"other", b""]
    for i, p in enumerate(small_vectors()):
        nm = names[i % len(names)]
        vectors.append({"name_hex": nm.hex(), "payload_hex": p.hex(),
                        "expect": S.route_one(TOOLS, DEFAULT_CLASS, PROVIDERS, nm, p)})
    for i, (kind, n) in enumerate([("pattern", 2048), ("text", 2048), ("text", 16383), ("text", 16384),
                                   ("pattern", 16385), ("text", 16400), ("pattern", 32768 + 5),
                                   ("text", 3 * 16384), ("text", 70001)]):
        p = pattern(n) if kind == "pattern" else text(n)
        nm = names[i % len(names)]
        vectors.append({"name_hex": nm.hex(), "payload_gen": [kind, n],
                        "expect": S.route_one(TOOLS, DEFAULT_CLASS, PROVIDERS, nm, p)})

    # the one request the reference's dev client sends — cmd/dev_client/main.go:247-258.
    # tool_name "example_tool"; Parameters {"example_param":"value","timestamp":<unix>} rendered
    # as canonical JSON with a fixed timestamp (the Go code uses time.Now()).
    dev_payload = b'{"example_param":"value","timestamp":1700000000}'
    dev = {"name_hex": b"example_tool".hex(), "payload_hex": dev_payload.hex(),
           "secret_id": "secret-123",
           "metadata": {"client_version": "1.0.0", "request_source": "dev_client"},
           "expect": S.route_one(TOOLS, DEFAULT_CLASS, PROVIDERS, b"example_tool", dev_payload)}

    # mock.go:24-63 response literals (hand-written)
    responses = {
        "status_code": 200, "status_message": "Tool executed successfully",
        "example_tool": {"kind": "string", "prefix": "Mock execution of example_tool at "},
        "struct_tool": {"kind": "struct", "result": "success", "processed": True, "count": 42},
        "file_tool": {"kind": "file", "file_name": "example.txt", "mime_type": "text/plain",
                      "content": "This is mock file content"},
        "default": {"kind": "string", "format": "Unknown tool: %s"},
    }
    return {"tools": [[n.decode(), c] for n, c in TOOLS], "default_class": DEFAULT_CLASS,
            "providers": PROVIDERS, "tool_class": rows, "vectors": vectors, "dev_client_request": dev,
            "responses": responses}


if __name__ == "__main__":
    out = os.path.join(HERE, "route_golden.json")
    with open(out, "w") as f:
        json.dump(build(), f, indent=1, sort_keys=True)
    print("wrote", out, os.path.getsize(out), "bytes")
