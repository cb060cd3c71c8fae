// svmTrainMain.hpp -- training parameters shared between the driver and SvmTrain.
// Field for field the reference's state_model (reference svmTrainMain.hpp:4-19): same order, same
// types, so a driver written against the reference header compiles unchanged.
#ifndef SVMTRAIN_HPP
#define SVMTRAIN_HPP

typedef struct {

	int num_attributes;        // -a / --num-att
	int num_train_data;        // -x / --num-ex   (GLOBAL row count)
	float c;                   // -c / --cost      (default 1)
	float gamma;               // -g / --gamma     (reference default 1/num_attributes in INTEGER arithmetic == 0)
	float epsilon;             // -e / --epsilon   (default 0.001)
	char input_file_name[200]; // -f / --file-path
	char model_file_name[200]; // -m / --model
	int max_iter;              // -n / --max-iter  (default 150000)
	int cache_size;            // -s / --cache-size, number of kernel-row cache lines (default 10)

} state_model;

// defined by the driver (reference svmTrainMain.cpp:18); READ by SvmTrain, as in the reference
extern state_model state;
#endif
