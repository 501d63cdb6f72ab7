// pybind_module.cpp — Python module `gaussianblur` with the reference's surface
// (/root/reference/python/bindings/gaussianblur_py.cpp:7-32): classes ImgGeom{rows, cols,
// channels} and Image{data, geom}, function gaussianblur(image, sigma, apply_to_alpha) that blurs
// the held image in place. Additive: gaussianblur_array() takes any writable C-contiguous uint8
// buffer (numpy array) and blurs it in place without the list <-> vector copies.
#include <gaussianblur/gaussianblur.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <stdexcept>

namespace py = pybind11;

PYBIND11_MODULE(gaussianblur, m) {
  m.doc() = "B200-native LibGaussianBlur: FFT Gaussian blur of 8-bit images on the GPU.";

  py::class_<ImgGeom>(m, "ImgGeom", "Image geometry: rows, cols, channels.")
      .def(py::init<>())
      .def_readwrite("rows", &ImgGeom::rows)
      .def_readwrite("cols", &ImgGeom::cols)
      .def_readwrite("channels", &ImgGeom::channels);

  py::class_<Image>(m, "Image", "Interleaved 8-bit image: data (list of ints) and geom.")
      .def(py::init<>())
      .def_readwrite("data", &Image::data)
      .def_readwrite("geom", &Image::geom);

  m.def(
      "gaussianblur",
      [](Image &image, float sigma, bool apply_to_alpha) { gaussianblur::gaussianblur(image, sigma, apply_to_alpha); },
      py::arg("image"), py::arg("sigma"), py::arg("apply_to_alpha"),
      "Blur the image in place. sigma: Gaussian standard deviation; apply_to_alpha: also blur the 4th channel.");

  m.def(
      "gaussianblur_array",
      [](py::buffer buf, float sigma, bool apply_to_alpha) {
        py::buffer_info info = buf.request(true);
        if (info.itemsize != 1) throw std::invalid_argument("expected a uint8 buffer");
        if (info.ndim != 2 && info.ndim != 3) throw std::invalid_argument("expected an HW or HWC array");
        ImgGeom g{(int)info.shape[0], (int)info.shape[1], info.ndim == 3 ? (int)info.shape[2] : 1};
        py::ssize_t expect = 1;
        for (int d = info.ndim - 1; d >= 0; --d) {
          if (info.strides[d] != expect) throw std::invalid_argument("array must be C-contiguous");
          expect *= info.shape[d];
        }
        uint8_t *ptr = static_cast<uint8_t *>(info.ptr);
        py::gil_scoped_release release;
        gaussianblur::gaussianblur_buffer(ptr, g, sigma, apply_to_alpha);
      },
      py::arg("array"), py::arg("sigma"), py::arg("apply_to_alpha") = false,
      "Blur a writable C-contiguous uint8 HW / HWC array in place (channels 1, 3 or 4).");

  m.def("set_device", &gaussianblur::set_device, py::arg("device"));
}
