/*
 * example_inprocess.cpp -- what a Dash application looks like with the B200
 * evaluator linked in (compare dash/example/main.cpp:78-86 of the reference).
 * Built only where the reference sources exist:
 *   nvcc -x cu -std=c++17 -I$DASH/dash/include -Iinclude -Iintegration \
 *        integration/example_inprocess.cpp -Ldash_b200 -ldash_b200 -o example
 * tests/test_integration_compile.py compiles it (no link, no run) to keep the
 * adapter honest against the unmodified reference headers.
 */
#include <iostream>
#include <vector>

#include "circuit/circuit.h"
#include "circuit/layer/dense.h"
#include "circuit/layer/relu.h"
#include "garbling/garbled_circuit.h"

#include "dash_b200.h"
#include "dash_b200_gcm.h"
#include "dash_adapter.h"

int main() {
    // a Dense -> ReLU -> Dense circuit as in dash/test/test_relu.h:40-55
    auto circuit = new Circuit{new Dense{ScalarTensor<wandb_t>{{1, 0, 1, 0}, dim_t{2, 2}}, ScalarTensor<wandb_t>{{0, 0}, dim_t{2}}, -1, QuantizationMethod::SimpleQuant, 1},
                               new Relu{dim_t{2}},
                               new Dense{ScalarTensor<wandb_t>{{1, 0, 1, 0}, dim_t{2, 2}}, ScalarTensor<wandb_t>{{0, 0}, dim_t{2}}, -1, QuantizationMethod::SimpleQuant, 1}};
    auto gc = new GarbledCircuit(circuit, 8, 100.0f);

    if (dashb_init(0) != DASHB_OK) {  // was: init_cuda()
        std::cerr << dashb_last_error() << std::endl;
        return 1;
    }
    dashb_circuit* dev = nullptr;
    if (dash_b200_move(gc, &dev, /*max_batch=*/1) != DASHB_OK) {  // was: gc->cuda_move()
        std::cerr << dashb_last_error() << std::endl;
        return 1;
    }
    ScalarTensor<q_val_t> inputs{{3, -4}, dim_t{2}};
    auto g_inputs = gc->garble_inputs(inputs);
    std::vector<const int16_t*> in;
    for (auto* l : *g_inputs) in.push_back(l->get_components());
    std::vector<LabelTensor*> g_outputs;
    std::vector<int16_t*> out;
    for (auto q : gc->get_crt_base()) {
        g_outputs.push_back(new LabelTensor(q, circuit->get_output_dims()));
        out.push_back(g_outputs.back()->get_components());
    }
    // was: cuda_move_inputs + cuda_evaluate + cuda_move_outputs
    if (dashb_evaluate(dev, in.data(), 1, out.data()) != DASHB_OK) {
        std::cerr << dashb_last_error() << std::endl;
        return 1;
    }
    auto outputs = gc->decode_outputs(&g_outputs);
    // same values from the reference's own CPU evaluator
    auto cpu = gc->decode_outputs(gc->cpu_evaluate(g_inputs));
    std::cout << "b200:";
    for (size_t i = 0; i < outputs.size(); ++i) std::cout << ' ' << outputs.data()[i];
    std::cout << "  cpu:";
    for (size_t i = 0; i < cpu.size(); ++i) std::cout << ' ' << cpu.data()[i];
    std::cout << std::endl;
    dashb_circuit_destroy(dev);
    return 0;
}
