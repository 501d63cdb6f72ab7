// Desktop front-end: gaussianblur <smoothing_factor> <input_file> [apply_to_alpha]
//
// Same command line, messages and output naming (<name>_pffft.<ext>) as the reference's desktop example
// (/root/reference/examples/desktop/main.cpp:52-99), on top of the B200 library. The reference decodes
// through stb_image; this build carries no third-party codec, so the formats are the ones small enough
// to read and write by hand: BMP (uncompressed 24/32-bit), binary PPM (P6) and PAM (P7, RGB / RGB_ALPHA).
#include <gaussianblur/gaussianblur.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

std::string lower_ext(const std::string& file) {
  const size_t dot = file.find_last_of('.');
  std::string e = dot == std::string::npos ? "" : file.substr(dot + 1);
  for (char& c : e) c = (char)tolower((unsigned char)c);
  return e;
}

std::vector<uint8_t> slurp(const std::string& file) {
  std::ifstream in(file, std::ios::binary);
  if (!in) throw std::runtime_error("Failed to load image: " + file);
  return std::vector<uint8_t>((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
}

uint32_t le32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
uint16_t le16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
void put32(std::vector<uint8_t>& v, uint32_t x) {
  for (int i = 0; i < 4; ++i) v.push_back((uint8_t)(x >> (8 * i)));
}
void put16(std::vector<uint8_t>& v, uint16_t x) {
  v.push_back((uint8_t)x);
  v.push_back((uint8_t)(x >> 8));
}

// ---- BMP: BITMAPFILEHEADER + BITMAPINFOHEADER (or V4/V5), BI_RGB / BI_BITFIELDS with the usual BGRA masks ----
Image read_bmp(const std::vector<uint8_t>& f) {
  if (f.size() < 54 || f[0] != 'B' || f[1] != 'M') throw std::runtime_error("not a BMP file");
  const uint32_t data_off = le32(&f[10]), hdr = le32(&f[14]);
  const int32_t w = (int32_t)le32(&f[18]), hs = (int32_t)le32(&f[22]);
  const uint16_t bpp = le16(&f[28]);
  const uint32_t comp = le32(&f[30]);
  if (hdr < 40 || w <= 0 || hs == 0) throw std::runtime_error("unsupported BMP header");
  if ((bpp != 24 && bpp != 32) || (comp != 0 && !(comp == 3 && bpp == 32)))
    throw std::runtime_error("Image format not supported, only uncompressed 24/32-bit BMP");
  const int h = hs < 0 ? -hs : hs, ch = bpp / 8;
  const size_t stride = ((size_t)w * ch + 3) & ~(size_t)3;
  if (f.size() < data_off + stride * h) throw std::runtime_error("truncated BMP file");
  Image img;
  img.geom = ImgGeom{h, w, ch};
  img.data.resize((size_t)h * w * ch);
  for (int y = 0; y < h; ++y) {
    const uint8_t* src = &f[data_off + stride * (hs < 0 ? y : h - 1 - y)];
    uint8_t* dst = &img.data[(size_t)y * w * ch];
    for (int x = 0; x < w; ++x) {  // BGR(A) -> RGB(A)
      dst[x * ch + 0] = src[x * ch + 2];
      dst[x * ch + 1] = src[x * ch + 1];
      dst[x * ch + 2] = src[x * ch + 0];
      if (ch == 4) dst[x * ch + 3] = src[x * ch + 3];
    }
  }
  return img;
}

void write_bmp(const std::string& file, const Image& img) {
  const int w = img.geom.cols, h = img.geom.rows, ch = img.geom.channels;
  const size_t stride = ((size_t)w * ch + 3) & ~(size_t)3;
  const uint32_t hdr = ch == 4 ? 108 : 40;  // V4 header carries the alpha mask
  std::vector<uint8_t> o;
  o.reserve(14 + hdr + stride * h);
  o.push_back('B');
  o.push_back('M');
  put32(o, (uint32_t)(14 + hdr + stride * h));
  put32(o, 0);
  put32(o, 14 + hdr);
  put32(o, hdr);
  put32(o, (uint32_t)w);
  put32(o, (uint32_t)h);  // bottom-up
  put16(o, 1);
  put16(o, (uint16_t)(ch * 8));
  put32(o, ch == 4 ? 3 : 0);
  put32(o, (uint32_t)(stride * h));
  put32(o, 2835);
  put32(o, 2835);
  put32(o, 0);
  put32(o, 0);
  if (ch == 4) {
    put32(o, 0x00FF0000u);
    put32(o, 0x0000FF00u);
    put32(o, 0x000000FFu);
    put32(o, 0xFF000000u);
    put32(o, 0x73524742u);  // 'sRGB'
    for (int i = 0; i < 12; ++i) put32(o, 0);
  }
  std::vector<uint8_t> row(stride, 0);
  for (int y = h - 1; y >= 0; --y) {
    const uint8_t* src = &img.data[(size_t)y * w * ch];
    for (int x = 0; x < w; ++x) {
      row[x * ch + 0] = src[x * ch + 2];
      row[x * ch + 1] = src[x * ch + 1];
      row[x * ch + 2] = src[x * ch + 0];
      if (ch == 4) row[x * ch + 3] = src[x * ch + 3];
    }
    o.insert(o.end(), row.begin(), row.end());
  }
  std::ofstream out(file, std::ios::binary);
  out.write(reinterpret_cast<const char*>(o.data()), (std::streamsize)o.size());
  if (!out) throw std::runtime_error("Failed to write image: " + file);
}

// ---- Netpbm: P6 (RGB) and P7 (TUPLTYPE RGB / RGB_ALPHA), maxval 255 ----
struct Tokens {
  const std::vector<uint8_t>& f;
  size_t pos;
  std::string next() {  // whitespace-separated token, '#' comments skipped
    for (;;) {
      while (pos < f.size() && isspace(f[pos])) ++pos;
      if (pos < f.size() && f[pos] == '#') {
        while (pos < f.size() && f[pos] != '\n') ++pos;
        continue;
      }
      break;
    }
    std::string t;
    while (pos < f.size() && !isspace(f[pos])) t.push_back((char)f[pos++]);
    return t;
  }
};

Image read_pnm(const std::vector<uint8_t>& f) {
  Tokens tk{f, 0};
  const std::string magic = tk.next();
  int w = 0, h = 0, ch = 0, maxval = 0;
  if (magic == "P6") {
    w = std::stoi(tk.next());
    h = std::stoi(tk.next());
    maxval = std::stoi(tk.next());
    ch = 3;
    ++tk.pos;  // the single whitespace byte after maxval
  } else if (magic == "P7") {
    for (;;) {
      const std::string k = tk.next();
      if (k == "ENDHDR" || k.empty()) break;
      const std::string v = tk.next();
      if (k == "WIDTH") w = std::stoi(v);
      if (k == "HEIGHT") h = std::stoi(v);
      if (k == "DEPTH") ch = std::stoi(v);
      if (k == "MAXVAL") maxval = std::stoi(v);
    }
    while (tk.pos < f.size() && f[tk.pos] != '\n') ++tk.pos;
    ++tk.pos;
  } else {
    throw std::runtime_error("not a binary PPM/PAM file");
  }
  if (w <= 0 || h <= 0 || maxval != 255) throw std::runtime_error("unsupported PPM/PAM header (8-bit samples only)");
  Image img;
  img.geom = ImgGeom{h, w, ch};
  const size_t n = (size_t)w * h * ch;
  if (f.size() < tk.pos + n) throw std::runtime_error("truncated PPM/PAM file");
  img.data.assign(f.begin() + (long)tk.pos, f.begin() + (long)(tk.pos + n));
  return img;
}

void write_pnm(const std::string& file, const Image& img, bool pam) {
  std::ostringstream hd;
  if (pam)
    hd << "P7\nWIDTH " << img.geom.cols << "\nHEIGHT " << img.geom.rows << "\nDEPTH " << img.geom.channels
       << "\nMAXVAL 255\nTUPLTYPE " << (img.geom.channels == 4 ? "RGB_ALPHA" : "RGB") << "\nENDHDR\n";
  else
    hd << "P6\n" << img.geom.cols << " " << img.geom.rows << "\n255\n";
  std::ofstream out(file, std::ios::binary);
  const std::string h = hd.str();
  out.write(h.data(), (std::streamsize)h.size());
  out.write(reinterpret_cast<const char*>(img.data.data()), (std::streamsize)img.data.size());
  if (!out) throw std::runtime_error("Failed to write image: " + file);
}

Image read_image(const std::string& file) {
  const std::string ext = lower_ext(file);
  const std::vector<uint8_t> f = slurp(file);
  Image img;
  if (ext == "bmp")
    img = read_bmp(f);
  else if (ext == "ppm" || ext == "pam")
    img = read_pnm(f);
  else
    throw std::runtime_error("Image format '" + ext + "' not supported by this build (no codec library): use bmp, ppm or pam");
  printf("Source image: %s %dx%d (%d)\n", file.c_str(), img.geom.cols, img.geom.rows, img.geom.channels);
  if (img.geom.channels != 3 && img.geom.channels != 4)
    throw std::runtime_error("Image format not supported, only RGB and RGBA images are supported");
  return img;
}

void write_image(const std::string& file, const Image& img) {
  const size_t dot = file.find_last_of('.');
  const std::string ext = lower_ext(file);
  const std::string out = file.substr(0, dot) + "_pffft." + file.substr(dot + 1);
  if (ext == "bmp")
    write_bmp(out, img);
  else
    write_pnm(out, img, ext == "pam" || img.geom.channels == 4);
}

void print_help() {
  std::cout << "Usage: gaussianblur <smoothing_factor> <input_file> [apply_to_alpha]\n";
  std::cout << "  <smoothing_factor> : The smoothing factor for the Gaussian blur (must be > 0).\n";
  std::cout << "  <input_file>       : The input image file (bmp, ppm, pam).\n";
  std::cout << "  [apply_to_alpha]   : Optional. If set to 1, the convolution is done on the 4th channel (alpha channel).\n";
  std::cout << "                       If not provided or set to 0, the convolution is done on the first 3 channels only.\n";
}

}  // namespace

int main(int argc, char* argv[]) {
  if (argc == 2 && (std::string(argv[1]) == "--help" || std::string(argv[1]) == "-h")) {
    print_help();
    return 0;
  }
  if (argc < 3 || argc > 4) {
    std::cerr << "Invalid number of arguments.\n";
    print_help();
    return 1;
  }
  try {
    const bool apply_to_alpha = argc == 4 && std::stoi(argv[3]) == 1;
    const std::string file_name(argv[2]);
    const float sigma = std::stof(argv[1]);
    if (sigma <= 0) {
      printf("Invalid smoothing factor\n");
      return 1;
    }
    Image image = read_image(file_name);
    gaussianblur::gaussianblur(image, sigma, apply_to_alpha);
    write_image(file_name, image);
  } catch (const std::exception& e) {
    std::cerr << e.what() << "\n";
    return 1;
  }
  return 0;
}
