"""Corner cases of the `.obj` grammar (scene_reader.py:29-104; vector.py:28-72, 143-173), shared by the golden
generator (tests/golden/make_golden.py runs the REFERENCE parser on each and records its result or exception class in
tests/golden/obj_corners.json) and by tests/test_obj_native.py (both product parsers must do the same)."""

GRAMMAR3 = """# header comment
   v 0 0 0   # trailing comment
v\t1.5\t0\t-0.0
v 1e0 1.25E+1 .5
v -3.141592653589793 2.718281828459045e-3 1e-7
v 0.1 0.2 0.30000000000000004
v 123456789012345678901234567890 1e308 4.9e-324
v inf -inf 1e400
usemtl something red
f 1 2 3
f -1 -2 -3
f 1 2 3 4 5
\t# tab then comment
f 0 1 2

g group one
s off
f +1 -6 3
"""

TRI = "v 0 0 0\nv 1 0 0\nv 0 1 0\n"

CASES = {
    "grammar_all_3d": GRAMMAR3,
    # a 4-coordinate vertex used together with 3-coordinate ones: V.__add__ refuses (vector.py:63-72)
    "grammar_mixed_dimension": GRAMMAR3.replace("v 1e0 1.25E+1 .5\n", "v 1e0 1.25E+1 .5 9.0\n"),
    "all_vertices_4d": "v 0 0 0 1\nv 1 0 0 1\nv 0 1 0 2\nf 1 2 3\n",              # cross_prod reads [0..2] only
    "all_vertices_5d": "v 0 0 0 1 7\nv 1 0 0 1 7\nv 0 1 0 2 7\nf 1 2 3\n",
    "mixed_dimension_unused_vertex": TRI + "v 5 5 5 5\nv 1 2\nf 1 2 3\n",       # never touched by a face: no error
    "mixed_dimension_second_vertex": "v 0 0 0\nv 1 0 0 1\nv 0 1 0\nf 1 2 3\n",
    "mixed_dimension_third_vertex": "v 0 0 0\nv 1 0 0\nv 0 1 0 1\nf 1 2 3\n",
    "all_vertices_2d": "v 0 0\nv 1 0\nv 0 1\nf 1 2 3\n",                       # u[2] in cross_prod: IndexError
    "all_vertices_1d": "v 0\nv 1\nv 2\nf 1 2 3\n",
    "vertex_without_coordinates": "v\n",                                       # V([]): coords[0] IndexError
    "face_with_two_indices": TRI + "f 1 2\n",
    "face_without_indices": TRI + "f\n",
    "face_index_past_the_end": TRI + "f 1 2 9\n",
    "face_refers_to_a_later_vertex": "v 0 0 0\nv 1 0 0\nf 1 2 3\nv 0 1 0\n",
    "face_zero_wraps_to_the_last_vertex_so_far": TRI + "f 0 1 2\nv 9 9 9\nf 0 1 2\n",
    "face_negative_past_the_start": TRI + "f -9 1 2\n",
    "face_negative_wraps_twice": TRI + "f -4 1 2\n",                           # 3 + (-4) = -1 -> vertexes[-1]
    "polygon_fan": "v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0.5 1.5 0\nv 0 1 0\nf 1 2 3 4 5\n",
    "slashes": TRI + "f 1/1/1 2/2/2 3/3/3\n",
    "bad_float": "v 0 0 zero\n",
    "hex_float": "v 0 0 0x10\n",
    "float_face_index": TRI + "f 1 2 3.0\n",
    "underscore_number": "v 1_0 0 0\nv 0 0 0\nv 0 1 0\nf 1 2 3\n",             # float('1_0') == 10.0 in Python
    "nan_vertex": "v nan 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n",
    "blank_last_line": "v 0 0 0\n   ",
    "blank_middle_line": "v 0 0 0\n   \nv 1 0 0\nv 0 1 0\nf 1 2 3\n",
    "empty_file": "",
    "comments_only": "# nothing\n  # here\n",
    "crlf": "v 0 0 0\r\nv 1 0 0\r\nv 0 1 0\r\nf 1 2 3",
    "unknown_keywords": TRI + "vn 0 0 1\nvt 0 0\ng a b\nusemtl m\no name\nf 1 2 3\n",
    "keyword_prefix_is_not_a_keyword": "vv 0 0 0\n" + TRI + "ff 1 2 3\nf 1 2 3\n",
    "grammar_with_a_repeated_vertex_in_a_face": GRAMMAR3.replace("f +1 -6 3", "f +1 -7 3"),       # -7 -> vertex 0 again: zero area
    "zero_area_triangle": "v 0 0 0\nv 1 1 1\nv 2 2 2\nf 1 2 3\n",               # normalize: 1 / 0.0
    "duplicate_vertex_triangle": "v 0 0 0\nv 0 0 0\nv 1 0 0\nf 1 2 3\n",
    "plus_signs": "v +1 +2.5 -0\nv 0 0 0\nv 0 1 0\nf +1 +2 +3\n",
    "double_sign": "v ++1 0 0\n",
    "hash_inside_token": TRI + "f 1 2 3#comment\n",
    "tabs_and_spaces": "\tv 5 5 5\n" + TRI + "f 2 3 4\n \t f 1 2 3\n",          # tabs become spaces after the comment pass
}
