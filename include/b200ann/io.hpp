// File formats on either side of the hot path, same names and layouts as the reference:
//   CNNS::load_fvecs / loadGT / load_centroids   CNNS/src/utils/aux_util.cpp:8-31, :33-54, :56-84
//   fvecs: per row [i32 dim][f32 x dim];  ivecs / ground truth: per row [u32 k][u32 x k];
//   centroids.fvecs: [i32 n_clusters][i32 m][u32 dim] then n_clusters*(m+1) fvecs rows
//   (written by CNNS/tests/cluster_IVF_nndescent.cpp:116-265).
// The reference prints and exit(1)s on a missing file; these throw std::runtime_error with the same text.
// Writers are provided so that GPU-built data can be handed back to the reference's drivers.
#ifndef B200ANN_IO_HPP
#define B200ANN_IO_HPP
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace CNNS {

inline std::vector<float> load_fvecs(const std::string &filename, unsigned &num, unsigned &dim) {
  std::ifstream in(filename, std::ios::binary);
  if (!in.is_open()) throw std::runtime_error("Open file error: " + filename);
  in.read((char *)&dim, 4);
  in.seekg(0, std::ios::end);
  size_t fsize = (size_t)in.tellg();
  num = (unsigned)(fsize / ((size_t)(dim + 1) * 4));
  std::vector<float> data((size_t)num * dim);
  in.seekg(0, std::ios::beg);
  for (size_t i = 0; i < num; ++i) {
    in.seekg(4, std::ios::cur);
    in.read((char *)(data.data() + i * dim), (std::streamsize)dim * sizeof(float));
  }
  return data;
}

inline std::vector<std::vector<unsigned>> loadGT(const char *filename) {
  std::ifstream in(filename, std::ios::binary | std::ios::in);
  if (!in) throw std::runtime_error("Cannot open file: " + std::string(filename));
  std::vector<std::vector<unsigned>> results;
  while (in) {
    unsigned GK;
    in.read((char *)&GK, sizeof(unsigned));
    if (!in) break;
    std::vector<unsigned> result(GK);
    in.read((char *)result.data(), (std::streamsize)GK * sizeof(unsigned));
    if (!in) break;
    results.push_back(result);
  }
  return results;
}

inline std::vector<float> load_centroids(const std::string &filename, int &n_clusters, int &m, unsigned &dim) {
  std::ifstream in(filename, std::ios::binary);
  if (!in.is_open()) throw std::runtime_error("Open file error: " + filename);
  in.read((char *)&n_clusters, sizeof(n_clusters));
  in.read((char *)&m, sizeof(m));
  in.read((char *)&dim, sizeof(dim));
  size_t total_points = (size_t)n_clusters * (size_t)(m + 1);
  std::vector<float> centroids(total_points * dim);
  for (size_t i = 0; i < total_points; ++i) {
    unsigned point_dim;
    in.read((char *)&point_dim, sizeof(point_dim));
    if (point_dim != dim) throw std::runtime_error("Dimension mismatch in centroids file");
    in.read((char *)(centroids.data() + i * dim), (std::streamsize)dim * sizeof(float));
  }
  return centroids;
}

inline void save_fvecs(const std::string &filename, const float *data, size_t num, unsigned dim) {
  std::ofstream out(filename, std::ios::binary);
  if (!out) throw std::runtime_error("Open file error: " + filename);
  for (size_t i = 0; i < num; ++i) {
    out.write((const char *)&dim, 4);
    out.write((const char *)(data + i * dim), (std::streamsize)dim * sizeof(float));
  }
}

inline void save_ivecs(const std::string &filename, const unsigned *ids, size_t num, unsigned k) {
  std::ofstream out(filename, std::ios::binary);
  if (!out) throw std::runtime_error("Open file error: " + filename);
  for (size_t i = 0; i < num; ++i) {
    out.write((const char *)&k, 4);
    out.write((const char *)(ids + i * k), (std::streamsize)k * sizeof(unsigned));
  }
}

}  // namespace CNNS
#endif
