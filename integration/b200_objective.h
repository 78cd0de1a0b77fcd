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
// tree happens; the host gradient vectors are filled as well whenever something else may read them.
#ifndef LAMBDAGAP_B200_INTEGRATION_B200_OBJECTIVE_H_
#define LAMBDAGAP_B200_INTEGRATION_B200_OBJECTIVE_H_

#include <LightGBM/config.h>
#include <LightGBM/objective_function.h>

#include <functional>
#include <memory>
#include <string>

#include "lambdagap_b200.h"

namespace LightGBM {

// Where the device copy of a host gradient vector lives.  One entry per B200Objective; the learner asks with the pointers
// TreeLearner::Train receives.  An entry answers once per GetGradients call (Take clears `fresh`), so gradients the caller
// changed or produced some other way are never replaced by a stale device copy.
struct DeviceGradients {
  const score_t* host_gradients = nullptr;
  const score_t* host_hessians = nullptr;
  const float* dev_gradients = nullptr;
  const float* dev_hessians = nullptr;
  data_size_t num_data = 0;
  int device_id = 0;
  bool fresh = false;

  static void Publish(const DeviceGradients& entry);
  static void Withdraw(const score_t* host_gradients);
  // true and *out filled when a fresh device copy of exactly these vectors exists on `device_id`
  static bool Take(const score_t* host_gradients, const score_t* host_hessians, data_size_t num_data, int device_id,
                   DeviceGradients* out);
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
  void ReleaseDevice();

  std::unique_ptr<ObjectiveFunction> host_;
  std::string type_;
  Config config_;
  int device_id_ = 0;
  data_size_t num_data_ = 0;
  bool on_device_ = false;     // false: every call goes to host_ (see Init for the cases)
  bool fill_host_ = false;     // also copy the gradients into the caller's vectors
  gbl_objective* dev_ = nullptr;
  double* score_dev_ = nullptr;
  float* gradients_dev_ = nullptr;
  float* hessians_dev_ = nullptr;
  mutable const double* registered_score_ = nullptr;
  mutable const score_t* published_ = nullptr;
};

}  // namespace LightGBM

#endif  // LAMBDAGAP_B200_INTEGRATION_B200_OBJECTIVE_H_
