// integration/b200_objective.h -- the gradient step of the drop-in library on the B200.
//
// The reference selects its objective in ObjectiveFunction::CreateObjectiveFunction (src/objective/objective_function.cpp:
// 19-113; with USE_CUDA and device_type=cuda it returns CUDA objectives whose gradients stay on the device, :21-69).  The
// library built by integration/Makefile keeps that translation unit as it is, compiled under another entry-point name, and
// b200_objective.cpp provides the entry point: the reference's own objective is created exactly as before and, for
// device_type=cuda and the objectives the device implements (regression, binary, lambdarank), wrapped in B200Objective.
//
// B200Objective forwards every call to the wrapped reference objective except GetGradients, which uploads the score
// vector GBDT hands over, runs the device objective (lambdagap_b200/csrc/objective.cu, rank_objective.cu) and leaves the
// gradients in HBM.  B200TreeLearner::Train finds them through DeviceGradients (below) by the host pointers GBDT passes and
// trains from the device copies, so neither the 10.5 M exp() of the host objective nor the upload of two float vectors per
// tree happens; the host gradient vectors are filled as well whenever no learner has announced that it trains from the
// device copies.  With num_gpu=N every GPU computes the gradients of its own rows (regression, binary).
#ifndef LAMBDAGAP_B200_INTEGRATION_B200_OBJECTIVE_H_
#define LAMBDAGAP_B200_INTEGRATION_B200_OBJECTIVE_H_

#include <LightGBM/config.h>
#include <LightGBM/objective_function.h>

#include <functional>
#include <memory>
#include <string>

#include "lambdagap_b200.h"

namespace LightGBM {

// How the rows are spread over GPUs: contiguous parts, one per GPU.  The learner announces its layout when it has built
// its device data; the objective computes each part's gradients on that part's GPU.
struct DeviceLayout {
  static constexpr int kMaxParts = 16;
  data_size_t num_data = 0;
  int num_parts = 0;
  data_size_t begin[kMaxParts] = {};
  data_size_t count[kMaxParts] = {};
  int device[kMaxParts] = {};
  bool operator==(const DeviceLayout& o) const;
};

// Where the device copies of a host gradient vector live.  One entry per B200Objective; the learner asks with the
// pointers TreeLearner::Train receives.  An entry answers once per GetGradients call (Take clears `fresh`), so gradients
// the caller changed or produced some other way are never replaced by a stale device copy.
struct DeviceGradients {
  const score_t* host_gradients = nullptr;
  const score_t* host_hessians = nullptr;
  DeviceLayout layout;
  const float* dev_gradients[DeviceLayout::kMaxParts] = {};  // part p: rows begin[p] .. begin[p] + count[p] - 1
  const float* dev_hessians[DeviceLayout::kMaxParts] = {};
  // multiclass: the host vectors hold num_class blocks of num_data entries (class-major, multiclass_objective.hpp:84-86),
  // GBDT calls Train once per class with the block's address (gbdt.cpp:385-402); part p's device arrays hold num_class
  // blocks of count[p] entries.  One bit per class that has not been trained from yet.
  int num_class = 1;
  unsigned long long fresh = 0;
  bool host_filled = false;  // the host vectors hold the same values

  static void Publish(const DeviceGradients& entry);
  static void Withdraw(const score_t* host_gradients);
  // A fresh entry for exactly these vectors and this layout: *out filled, true.  A fresh entry that cannot be used (other
  // layout, `usable` false) is fine when the host vectors were filled too and fatal otherwise -- never a silent mix-up.
  static bool Take(const score_t* host_gradients, const score_t* host_hessians, const DeviceLayout& layout, bool usable,
                   DeviceGradients* out);
  // learners that will train from device copies
  static void AddConsumer(const DeviceLayout& layout);
  static void RemoveConsumer(const DeviceLayout& layout);
  // the learner created last over `num_data` rows on `num_parts` GPUs (one part: on GPU `device_id`)
  static bool FindConsumer(data_size_t num_data, int num_parts, int device_id, DeviceLayout* out);
};

class B200Objective : public ObjectiveFunction {
 public:
  // takes ownership of `host` (the reference's objective for `type`)
  B200Objective(ObjectiveFunction* host, const std::string& type, const Config& config);
  ~B200Objective() override;

  // the device implements these (and only without the options listed in Supported's body)
  static bool Supported(const std::string& type, const Config& config);

  void Init(const Metadata& metadata, data_size_t num_data) override;
  void GetGradients(const double* score, score_t* gradients, score_t* hessians) const override;
  void GetGradients(const double* score, const data_size_t num_sampled_queries, const data_size_t* sampled_query_indices,
                    score_t* gradients, score_t* hessians) const override;

  const char* GetName() const override { return host_->GetName(); }
  bool IsConstantHessian() const override { return host_->IsConstantHessian(); }
  bool IsRenewTreeOutput() const override { return host_->IsRenewTreeOutput(); }
  double RenewTreeOutput(double ori_output, std::function<double(const label_t*, int)> residual_getter,
                         const data_size_t* index_mapper, const data_size_t* bagging_mapper,
                         data_size_t num_data_in_leaf) const override {
    return host_->RenewTreeOutput(ori_output, residual_getter, index_mapper, bagging_mapper, num_data_in_leaf);
  }
  double BoostFromScore(int class_id) const override { return host_->BoostFromScore(class_id); }
  bool ClassNeedTrain(int class_id) const override { return host_->ClassNeedTrain(class_id); }
  bool SkipEmptyClass() const override { return host_->SkipEmptyClass(); }
  int NumModelPerIteration() const override { return host_->NumModelPerIteration(); }
  int NumPredictOneRow() const override { return host_->NumPredictOneRow(); }
  bool NeedAccuratePrediction() const override { return host_->NeedAccuratePrediction(); }
  data_size_t NumPositiveData() const override { return host_->NumPositiveData(); }
  void ConvertOutput(const double* input, double* output) const override { host_->ConvertOutput(input, output); }
  std::string ToString() const override { return host_->ToString(); }

 private:
  struct Part {
    gbl_objective* objective = nullptr;
    double* score = nullptr;
    float* gradients = nullptr;
    float* hessians = nullptr;
  };
  void ReleaseDevice() const;
  void EnsureDevice() const;  // first GetGradients: the learner exists by then and has announced its layout
  template <typename F>
  void ForEachPart(F fn) const;

  std::unique_ptr<ObjectiveFunction> host_;
  std::string type_;
  Config config_;
  data_size_t num_data_ = 0;
  const Metadata* metadata_ = nullptr;
  bool force_fill_host_ = false;
  // device state, built on first use
  mutable bool decided_ = false;
  mutable bool on_device_ = false;   // false: every call goes to host_ (see EnsureDevice for the cases)
  mutable bool fill_host_ = false;   // also copy the gradients into the caller's vectors
  int num_class_ = 1;
  mutable DeviceLayout layout_;
  mutable Part parts_[DeviceLayout::kMaxParts];
  mutable const double* registered_score_ = nullptr;
  mutable const score_t* published_ = nullptr;
};

}  // namespace LightGBM

#endif  // LAMBDAGAP_B200_INTEGRATION_B200_OBJECTIVE_H_
