// oracle/ref_tap.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// A few read-only taps into the UNMODIFIED reference CPU library, compiled against the reference's own
// headers where they lie (-I/root/reference/include) and linked to oracle/_ref/lib_lightgbm.so.  They let
// the parity tests see what the C API does not expose: the per-feature BinMapper (bin upper bounds,
// most-frequent / default bin, missing type; include/LightGBM/bin.h:86-260) and the mapper bin of every
// cell (BinIterator::Get, include/LightGBM/bin.h:262-290, src/io/dense_bin.hpp:24-50).
// A DatasetHandle of the reference C API is a plain `Dataset*` (src/c_api.cpp:1303-1405).
#include <LightGBM/bin.h>
#include <LightGBM/dataset.h>

#include <cstdint>
#include <memory>

using LightGBM::BinIterator;
using LightGBM::BinMapper;
using LightGBM::Dataset;

extern "C" {

int RefTap_NumFeatures(void* handle) {
  return reinterpret_cast<Dataset*>(handle)->num_features();
}

int RefTap_NumTotalFeatures(void* handle) {
  return reinterpret_cast<Dataset*>(handle)->num_total_features();
}

// Fills the scalars of used (inner) feature `inner`; `upper_bounds` must hold num_bin doubles (<= cap).
int RefTap_FeatureInfo(void* handle, int inner, int* real_index, int* num_bin, int* missing_type,
                       int* default_bin, int* most_freq_bin, double* upper_bounds, int cap) {
  auto* ds = reinterpret_cast<Dataset*>(handle);
  if (inner < 0 || inner >= ds->num_features()) return -1;
  const BinMapper* m = ds->FeatureBinMapper(inner);
  *real_index = ds->RealFeatureIndex(inner);
  *num_bin = m->num_bin();
  *missing_type = static_cast<int>(m->missing_type());
  *default_bin = static_cast<int>(m->GetDefaultBin());
  *most_freq_bin = static_cast<int>(m->GetMostFreqBin());
  if (m->num_bin() > cap) return -2;
  for (int b = 0; b < m->num_bin(); ++b) upper_bounds[b] = m->BinToValue(b);
  return 0;
}

// Mapper bin (0..num_bin-1, most-frequent bin restored) of every row for used feature `inner`.
int RefTap_FeatureBins(void* handle, int inner, uint8_t* out) {
  auto* ds = reinterpret_cast<Dataset*>(handle);
  if (inner < 0 || inner >= ds->num_features()) return -1;
  std::unique_ptr<BinIterator> it(ds->FeatureIterator(inner));
  it->Reset(0);
  const int n = ds->num_data();
  for (int i = 0; i < n; ++i) out[i] = static_cast<uint8_t>(it->Get(i));
  return 0;
}

}  // extern "C"
