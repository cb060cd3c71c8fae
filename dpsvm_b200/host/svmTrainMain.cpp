// svmTrainMain.cpp -- command-line driver: same flags, defaults and stdout lines as the reference's
// driver (reference svmTrainMain.cpp:46-136, 142-365), but the SMO loop itself runs on the device
// (SvmTrain::train) instead of as a host loop around train_step1 / Allgather / train_step2.
//
//   svmTrain -a <attrs> -x <rows> -f <csv> -m <model> [-c C] [-g gamma] [-e eps] [-n max_iter] [-s lines]
//   extra:   --seq-model   write seq.cpp's b-less model variant (readable by the reference's seq_test)
//            --host-loop   drive train_step1/train_step2 from the host exactly as the reference main does
#include <getopt.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <iostream>
#include <vector>

#include "dpsvm_cuda.h"
#include "dpsvm_model.hpp"
#include "parse.hpp"
#include "svmTrain.h"
#include "svmTrainMain.hpp"

state_model state;

using namespace std;

static bool g_seq_model = false;
static bool g_host_loop = false;

static void usage_exit()
{
    cerr << "   Command Line:\n\n"
            "   -a/--num-att        :  [REQUIRED] The number of attributes/features\n"
            "   -x/--num-ex         :  [REQUIRED] The number of training examples\n"
            "   -f/--file-path      :  [REQUIRED] Path to the training file\n"
            "   -m/--model          :  [REQUIRED] Path of model to be saved\n"
            "   -c/--cost           :  Parameter c of the SVM (default 1)\n"
            "   -g/--gamma          :  Parameter gamma of the radial basis function exp(-gamma*|u-v|^2)\n"
            "                          (default: 1/num-att in integer arithmetic, i.e. 0, as in the reference)\n"
            "   -e/--epsilon        :  Tolerance of termination criterion (default 0.001)\n"
            "   -n/--max-iter       :  Maximum number of iterations (default 150,000)\n"
            "   -s/--cache-size     :  Size of cache (num cache lines, default 10)\n"
            "      --seq-model      :  write the model without the b line (seq.cpp / seq_test format)\n"
            "      --host-loop      :  host-driven train_step1/train_step2 loop (reference main's structure)\n\n";
    exit(-1);
}

static void parse_arguments(int argc, char* argv[])
{
    static struct option opts[] = {
        {"num-att", required_argument, 0, 'a'},   {"num-ex", required_argument, 0, 'x'},
        {"cost", required_argument, 0, 'c'},      {"gamma", required_argument, 0, 'g'},
        {"file-path", required_argument, 0, 'f'}, {"epsilon", required_argument, 0, 'e'},
        {"max-iter", required_argument, 0, 'n'},  {"model", required_argument, 0, 'm'},
        {"cache-size", required_argument, 0, 's'}, {"seq-model", no_argument, 0, 1000},
        {"host-loop", no_argument, 0, 1001},      {0, 0, 0, 0}};
    // defaults: reference svmTrainMain.cpp:63-71
    state.epsilon = 0.001f;
    state.c = 1;
    state.num_attributes = -1;
    state.num_train_data = -1;
    state.gamma = -1;
    state.input_file_name[0] = 0;
    state.model_file_name[0] = 0;
    state.max_iter = 150000;
    state.cache_size = 10;
    for (;;) {
        int idx = 0;
        const int o = getopt_long(argc, argv, "a:x:c:g:f:e:n:m:s:", opts, &idx);
        if (o == -1) break;
        switch (o) {
            case 'a': state.num_attributes = atoi(optarg); break;
            case 'x': state.num_train_data = atoi(optarg); break;
            case 'c': state.c = atof(optarg); break;
            case 'g': state.gamma = atof(optarg); break;
            case 'f': snprintf(state.input_file_name, sizeof(state.input_file_name), "%s", optarg); break;
            case 'e': state.epsilon = atof(optarg); break;
            case 'n': state.max_iter = atoi(optarg); break;
            case 'm': snprintf(state.model_file_name, sizeof(state.model_file_name), "%s", optarg); break;
            case 's': state.cache_size = atoi(optarg); break;
            case 1000: g_seq_model = true; break;
            case 1001: g_host_loop = true; break;
            default:
                cerr << "\nERROR: Unknown option: -" << o << "\n";
                usage_exit();
        }
    }
    if (!state.input_file_name[0] || !state.model_file_name[0]) {
        cerr << "Enter a valid file name\n";
        usage_exit();
    }
    if (state.num_attributes <= 0 || state.num_train_data <= 0) {
        cerr << "Missing a required parameter, or invalid parameter\n";
        usage_exit();
    }
    // reference svmTrainMain.cpp:131-134: integer division, so the default gamma is 0 for any
    // num_attributes > 1.  Kept (nothing is fixed silently); pass -g.
    if (state.gamma < 0) state.gamma = 1 / state.num_attributes;
}

// The reference main's loop body (svmTrainMain.cpp:235-310) for one rank, over the drop-in API.
static int host_loop(SvmTrain& svm, std::vector<int>& raw_y, std::vector<float>& alpha, float& b_lo, float& b_hi)
{
    int num_iter = 0;
    do {
        svm.train_step1();
        int i_hi, i_lo;
        dpsvm_merge_candidates(svm.rv, 1, &i_hi, &i_lo, &b_hi, &b_lo);
        const float eta = svm.rbf_kernel(i_hi, i_hi) + svm.rbf_kernel(i_lo, i_lo) - (2 * svm.rbf_kernel(i_lo, i_hi));
        float a_hi, a_lo;
        dpsvm_alpha_update(eta, raw_y[i_hi], raw_y[i_lo], b_hi, b_lo, alpha[i_hi], alpha[i_lo], state.c, &a_hi, &a_lo);
        alpha[i_lo] = a_lo;
        alpha[i_hi] = a_hi;
        svm.train_step2(i_hi, i_lo, a_hi, a_lo);
        num_iter++;
    } while ((b_lo > (b_hi + (2 * state.epsilon))) && num_iter < state.max_iter);
    return num_iter;
}

int main(int argc, char* argv[])
{
    parse_arguments(argc, argv);

    std::vector<float> raw_x((size_t)state.num_train_data * state.num_attributes, 0);
    std::vector<int> raw_y(state.num_train_data, 0);
    populate_data(raw_x, raw_y, state.num_train_data, state.num_attributes, state.input_file_name);
    cout << "Populated Data from input file at node: " << 0 << '\n';
    cout << 0 << '\t' << state.num_train_data << '\n';            // shard table (svmTrainMain.cpp:185-189)

    SvmTrain svm(state.num_train_data, 0);
    svm.setup(raw_x, raw_y);
    cout << "SETUP DONE\n";

    const auto t0 = std::chrono::steady_clock::now();
    float b_lo = 0, b_hi = 0;
    int num_iter;
    std::vector<float> alpha;
    if (g_host_loop) {
        alpha.assign((size_t)state.num_train_data, 0.0f);
        num_iter = host_loop(svm, raw_y, alpha, b_lo, b_hi);
    } else {
        num_iter = svm.train();
        b_lo = svm.b_lo;
        b_hi = svm.b_hi;
        alpha = svm.alpha();
    }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    // the reference prints integer seconds (CycleTimer::currentSeconds truncated, svmTrainMain.cpp:312-314)
    cout << "TOTAL TIME TAKEN in seconds: " << (unsigned long long)secs << "\n";
    printf("Training time: %.6f s, %d iterations, %.1f iterations/s\n", secs, num_iter, num_iter / (secs > 0 ? secs : 1e-9));
    if (b_lo > (b_hi + (2 * state.epsilon))) {
        cout << "Could not converge in " << num_iter << " iterations. SVM training has been stopped\n";
    } else {
        cout << "Converged at iteration number: " << num_iter << "\n";
    }
    if (svm.cache_lookups() > 0)
        printf("Kernel-row cache: %lld hits / %lld lookups (%.1f %%)\n", svm.cache_hits(), svm.cache_lookups(),
               100.0 * svm.cache_hits() / svm.cache_lookups());

    svm.b = (b_lo + b_hi) / 2;                                      // svmTrainMain.cpp:329
    cout << "b: " << svm.b << "\n";
    svm.test_setup();
    const float train_accuracy = svm.get_train_accuracy();
    cout << "Training accuracy: " << train_accuracy << "\n";

    if (!dpsvm_write_model(state.model_file_name, raw_x.data(), raw_y.data(), alpha.data(), svm.b, state.gamma,
                           state.num_train_data, state.num_attributes, !g_seq_model)) {
        cout << "Model output file " << state.model_file_name << " could not be opened for writing.\n";
        exit(-1);
    }
    cout << "Training model has been saved to the file " << state.model_file_name << "\n";
    return 0;
}
