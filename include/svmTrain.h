// svmTrain.h -- the SMO trainer object, source-compatible with the reference's class SvmTrain
// (reference svmTrain.h:48-140) for everything its driver calls (svmTrainMain.cpp:194-338).
//
// Differences, all forced by "no Thrust / cuBLAS in the product":
//   * private members are behind a pimpl; device state lives in libdpsvm.so (include/dpsvm_cuda.h);
//   * lookup_cache returns a raw device pointer instead of thrust::device_vector<float>&
//     (only SvmTrain itself calls it in the reference, svmTrain.cu:203,238);
//   * additions: train() (the whole loop of svmTrainMain.cpp:235-310 on the device), accessors.
// Errors keep the reference's behaviour: message on stdout, exit(EXIT_FAILURE)
// (svmTrain.cu:166-179, 586-590).
#ifndef SVMTRAIN
#define SVMTRAIN

#include <vector>

typedef struct {

	int I_hi;
	int I_lo;
	float b_hi;
	float b_lo;

} step1_rv;

class SvmTrain {

	private:

		struct Impl;
		Impl* impl;

		int num_train_data;   // rows of this shard            (reference svmTrain.h:79)
		int start;            // first global row of the shard (:81)
		int end;

	public:

		float* rv;            // {I_hi, I_lo, b_hi, b_lo}, indices as floats (reference svmTrain.h:108, svmTrain.cu:478-481)
		float b;

		SvmTrain(int n_data, int d);     // n_data shard rows, d = global index of the first one (svmTrain.cu:305-316)
		~SvmTrain();

		void setup(std::vector<float>& raw_x, std::vector<int>& raw_y);   // svmTrain.cu:319-395

		void train_step1();                                               // svmTrain.cu:469-483

		void train_step2(int I_hi, int I_lo, float alpha_hi, float alpha_lo);   // svmTrain.cu:485-497

		void init_cuda_handles();        // kept for source compatibility (streams live in the context)

		void destroy_cuda_handles();

		int update_f(int I_lo, int I_hi, int y_lo, int y_hi, float alpha_lo_old, float alpha_hi_old, float alpha_lo_new, float alpha_hi_new);   // svmTrain.cu:190-302

		float* lookup_cache(int I_idx, bool& cache_hit);                  // svmTrain.cu:142-156

		float clip_value(float num, float low, float high);               // svmTrain.cu:674-682

		float rbf_kernel(int i1, int i2);                                 // HOST kernel on the host copy of X, svmTrain.cu:696-714

		void get_x(float* x, float* x_copy, int idx, int num_attributes); // svmTrain.cu:684-693

		float get_train_accuracy();                                       // svmTrain.cu:633-665

		void test_setup();                                                // svmTrain.cu:569-592

		void aggregate_sv();                                              // svmTrain.cu:595-631

		void destroy_t_cuda_handles();

		// ---- additions -------------------------------------------------------------------
		// Runs the loop of svmTrainMain.cpp:235-310 on the device until convergence / state.max_iter
		// (or max_new_iter more iterations when > 0).  Returns the total iteration count; fills b_hi,
		// b_lo, converged and the host alpha copy (the driver-side alpha[] of svmTrainMain.cpp:223).
		int train(int max_new_iter = 0);
		float b_hi, b_lo;
		bool converged;
		const std::vector<float>& alpha() const;     // global-length host copy, refreshed by train()/test_setup()
		void set_alpha(const float* alpha_full);     // for drivers that keep alpha themselves (reference main)
		long long cache_hits() const;
		long long cache_lookups() const;
		int num_sv() const;                          // set by aggregate_sv()
};

#endif
