// reads the files tests/test_io_formats.py wrote with the reference-named loaders (include/b200ann/io.hpp),
// checks them against values passed on the command line, writes them back with the writers.
#include <b200ann/io.hpp>
#include <cstdio>
#include <cstdlib>
#define CHECK(c) do { if (!(c)) { printf("CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)
int main(int argc, char **argv) {
  std::string dir = argv[1];
  unsigned num = 0, dim = 0;
  std::vector<float> x = CNNS::load_fvecs(dir + "/a.fvecs", num, dim);
  CHECK(num == (unsigned)atoi(argv[2]) && dim == (unsigned)atoi(argv[3]));
  CHECK(x[0] == 0.5f && x[(size_t)num * dim - 1] == (float)atof(argv[4]));
  auto gt = CNNS::loadGT((dir + "/gt.ivecs").c_str());
  CHECK(gt.size() == 7 && gt[0].size() == 5 && gt[6][4] == 6 * 5 + 4);
  int nc = 0, m = 0; unsigned cd = 0;
  std::vector<float> c = CNNS::load_centroids(dir + "/centroids.fvecs", nc, m, cd);
  CHECK(nc == 3 && m == 2 && cd == dim && c.size() == (size_t)9 * dim && c[dim] == x[dim]);
  CNNS::save_fvecs(dir + "/b.fvecs", x.data(), num, dim);
  std::vector<unsigned> flat;
  for (auto &r : gt) flat.insert(flat.end(), r.begin(), r.end());
  CNNS::save_ivecs(dir + "/gt2.ivecs", flat.data(), gt.size(), 5);
  bool threw = false;
  try { CNNS::load_fvecs(dir + "/missing.fvecs", num, dim); } catch (std::runtime_error &) { threw = true; }
  CHECK(threw);
  printf("io ok\n");
  return 0;
}
