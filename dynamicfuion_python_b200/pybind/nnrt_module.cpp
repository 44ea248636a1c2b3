// PYBIND11_MODULE(nnrt, m): the reference's module name and submodule tree (cpp/pybind/nnrt_pybind.cpp:23-40), with the hot path's host
// classes in C++ over libnnrt_b200.so. `nnrt.alignment` is new (the reference never bound DeformableMeshToImageFitter).
#include "bridge.h"

PYBIND11_MODULE(nnrt, m) {
	m.doc() = "B200-native dense-depth non-rigid tracking path behind the nnrt API";
	m.attr("_GLIBCXX_USE_CXX11_ABI") = py::bool_(_GLIBCXX_USE_CXX11_ABI != 0);
	m.attr("b200_c_abi_version") = nnrt_b200_version();
	py::module_ core = m.def_submodule("core", "Tensor-based core module.");
	nnrt::bind_core(core);
	py::module_ geometry = m.def_submodule("geometry", "Warp fields and geometry functions.");
	nnrt::bind_geometry(geometry);
	py::module_ rendering = m.def_submodule("rendering", "Rasterization.");
	nnrt::bind_rendering(rendering);
	nnrt::bind_shaders(rendering);
	py::module_ alignment = m.def_submodule("alignment", "Dense-depth mesh-to-image fitter.");
	nnrt::bind_alignment(alignment);
	py::module_ alignment_functional = alignment.def_submodule("functional", "Un-fused stages of the data term.");
	nnrt::bind_alignment_functional(alignment_functional);
	py::module_ io = m.def_submodule("io", "Tensor files in the reference's binary format.");
	nnrt::bind_io(io);
}
