"""Known answers copied from the Apache Avro 1.x specification, section "Binary Encoding" — the only byte-level Avro
examples that do not come from this repository's own encoder.  Shared by tests/test_avro_spec_cpu.py (oracle + Python
encoder) and tests/test_decode_gpu.py / test_avro_encode_gpu.py (CUDA decoder + encoder).

  * "Primitive Types": int and long are written using variable-length zig-zag coding; the spec's table:
        value  0   -1   1   -2   2   ...  -64   64
        hex    00  01   02  03   04  ...  7f    80 01
  * a string is a long byte count followed by that many bytes of UTF-8: "foo" → 06 66 6f 6f
  * float / double: 4 / 8 bytes, IEEE-754 bits (Java floatToIntBits / doubleToLongBits) little-endian
  * boolean: a single byte 0 or 1;  null: zero bytes
  * "Complex Types / Records": fields in declaration order, nothing else.  The spec's example — schema
        {"type": "record", "name": "test", "fields": [{"name": "a", "type": "long"}, {"name": "b", "type": "string"}]}
    with a = 27, b = "foo" encodes as 36 06 66 6f 6f
  * "Unions": a long with the zero-based branch position, then the value.  ["null","string"]: null → 00, "a" → 02 02 61
"""
import struct

ZIGZAG_VALUES = [0, -1, 1, -2, 2, -64, 64]
ZIGZAG_HEX = "00 01 02 03 04 7f 80 01"

SCHEMA_ZIGZAG_LONGS = {"type": "record", "name": "z", "fields": [{"name": f"v{i}", "type": "long"} for i in range(7)]}
SCHEMA_ZIGZAG_INTS = {"type": "record", "name": "z", "fields": [{"name": f"v{i}", "type": "int"} for i in range(7)]}

SCHEMA_SPEC_RECORD = {"type": "record", "name": "test", "fields": [{"name": "a", "type": "long"}, {"name": "b", "type": "string"}]}
SPEC_RECORD_HEX = "36 06 66 6f 6f"          # a = 27, b = "foo"

SCHEMA_UNION_STRING = {"type": "record", "name": "u", "fields": [{"name": "s", "type": ["null", "string"]}]}
UNION_NULL_HEX = "00"
UNION_A_HEX = "02 02 61"                    # branch 1, length 1, "a"
UNION_TRUE_HEX = "02 08 74 72 75 65"        # branch 1, length 4, "true"  (the FAILURE_OCCURRED label of the KSQL schema)
UNION_FALSE_HEX = "02 0a 66 61 6c 73 65"

SCHEMA_FLOATS = {"type": "record", "name": "f", "fields": [{"name": "f", "type": "float"}, {"name": "d", "type": "double"},
                                                            {"name": "b", "type": "boolean"}, {"name": "n", "type": "null"},
                                                            {"name": "g", "type": "float"}]}
# floatToIntBits(1.0f) = 0x3f800000, doubleToLongBits(-2.5) = 0xc004000000000000, true, (null), floatToIntBits(-0.15625f) = 0xbe200000
FLOATS_VALUES = [1.0, -2.5, 1.0, -0.15625]
FLOATS_HEX = "00 00 80 3f" " 00 00 00 00 00 00 04 c0" " 01" " 00 00 20 be"
assert bytes.fromhex(FLOATS_HEX) == struct.pack("<f", 1.0) + struct.pack("<d", -2.5) + b"\x01" + struct.pack("<f", -0.15625)

SCHEMA_UNION_NUMBERS = {"type": "record", "name": "n", "fields": [{"name": "d", "type": ["null", "double"]},
                                                                   {"name": "i", "type": ["null", "int"]},
                                                                   {"name": "e", "type": ["double", "null"]}]}
# branch 1 + 0.5 ; branch 1 + int 300 (zig-zag 600 = d8 04) ; [double, null] with null = branch 1
UNION_NUMBERS_HEX = "02 00 00 00 00 00 00 e0 3f" " 02 d8 04" " 02"
UNION_NUMBERS_VALUES = [0.5, 300.0, 0.0]


def h(s: str) -> bytes:
    return bytes.fromhex(s.replace(" ", ""))
