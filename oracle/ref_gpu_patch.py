"""ORACLE / BASELINE build helper (test infrastructure only).

Writes a scratch copy of the reference's src/testbed_geometry_training.cu in which the ONE signature that does not
match its declaration is completed: Testbed::render_geometry_mesh is defined with 9 parameters
(src/testbed_geometry_training.cu:2202-2212) but declared -- and called, src/testbed.cu:4852-4869 -- with 14
(include/neural-graphics-primitives/testbed.h:530-545). The five missing parameters are appended, unused, exactly as
the header spells them. Nothing else in the file changes. The output lives under oracle/_ref/obj/ (git-ignored)."""
import sys

src, dst = sys.argv[1], sys.argv[2]
text = open(src, encoding="utf-8", errors="surrogateescape").read()
head = "void Testbed::render_geometry_mesh("
i = text.index(head)
needle = "\tint visualized_dimension\n) {"
j = text.index(needle, i)
assert j - i < 600, "render_geometry_mesh signature not where expected"
extra = (
    "\tint visualized_dimension,\n"
    "\tCudaDevice& device,\n"
    "\tconst std::shared_ptr<NerfNetwork<network_precision_t>>& nerf_network,\n"
    "\tconst uint8_t* density_grid_bitfield,\n"
    "\tconst mat4x3& camera_matrix1,\n"
    "\tconst vec4& rolling_shutter\n) {"
)
text = text[:j] + extra + text[j + len(needle):]
open(dst, "w", encoding="utf-8", errors="surrogateescape").write(text)
