// Stand-alone driver of the native FC parameter server (counterpart of MPI_code/src/distributed_nn.cpp and
// single_machine_nn.cpp). Build: `make -C csrc distributed`; run: `make -C csrc distributed_run`.
//   distributed_nn_native [--workers 6] [--collect 2] [--iters 100] [--batch 128] [--lr 0.001] [--data DIR]
//                         [--delay-mask 0x7 --delay-us 500] [--no-shortcircuit] [--timeline] [--selftest] [--host-protocol]
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/pdnn_b200.h"

int main(int argc, char** argv) {
    pdnn_fc_ps_config_t c;
    memset(&c, 0, sizeof(c));
    const int dims[] = {784, 500, 500, 800, 800, 200, 100, 100, 10};    // MPI_code/src/distributed_nn.cpp:35-47
    c.n_dims = 9;
    memcpy(c.dims, dims, sizeof(dims));
    c.batch = 128; c.lr = 1e-3f; c.n_workers = 6; c.n_to_collect = 2; c.n_train_iters = 100;
    c.shortcircuit = 1; c.with_evaluator = 1; c.eval_batches = 4; c.seed = 1; c.synthetic_examples = 8192;
    c.device_protocol = 1;      // roles talk through the parameter-server kernels (device flags + peer stores)
    snprintf(c.out_dir, sizeof(c.out_dir), "outfiles");
    for (int i = 1; i < argc; ++i) {
        const char* a = argv[i];
        const char* v = i + 1 < argc ? argv[i + 1] : "";
        if (!strcmp(a, "--workers")) { c.n_workers = atoi(v); ++i; }
        else if (!strcmp(a, "--collect")) { c.n_to_collect = atoi(v); ++i; }
        else if (!strcmp(a, "--iters")) { c.n_train_iters = atoi(v); ++i; }
        else if (!strcmp(a, "--batch")) { c.batch = atoi(v); ++i; }
        else if (!strcmp(a, "--lr")) { c.lr = (float)atof(v); ++i; }
        else if (!strcmp(a, "--data")) { snprintf(c.data_dir, sizeof(c.data_dir), "%s", v); ++i; }
        else if (!strcmp(a, "--out")) { snprintf(c.out_dir, sizeof(c.out_dir), "%s", v); ++i; }
        else if (!strcmp(a, "--delay-mask")) { c.delay_worker_mask = strtoull(v, nullptr, 0); ++i; }
        else if (!strcmp(a, "--delay-us")) { c.delay_us_per_layer = atoll(v); ++i; }
        else if (!strcmp(a, "--no-shortcircuit")) c.shortcircuit = 0;
        else if (!strcmp(a, "--timeline")) c.generate_timeline = 1;
        else if (!strcmp(a, "--host-protocol")) c.device_protocol = 0;
        else if (!strcmp(a, "--help") || !strcmp(a, "-h")) {
            printf("usage: %s [--workers 6] [--collect 2] [--iters 100] [--batch 128] [--lr 0.001] [--data DIR] [--out DIR]\n"
                   "          [--delay-mask 0x7 --delay-us 500] [--no-shortcircuit] [--timeline] [--host-protocol] [--selftest]\n"
                   "FC parameter server; the roles are threads of this process: master on GPU 0, --workers workers round-robin over the\n"
                   "other visible GPUs (all on GPU 0 when there is only one), evaluator on the last GPU.\n"
                   "--collect k: gradients the master waits for per layer; --delay-mask / --delay-us: injected per-layer delay of the\n"
                   "workers in the mask; --host-protocol: host-thread transport instead of the device-flag protocol; --selftest: train\n"
                   "a small net on one GPU and report the loss (MNIST idx files are read from --data, synthetic data otherwise).\n", argv[0]);
            return 0;
        }
        else if (!strcmp(a, "--selftest")) {
            float l0, l1;
            int rc = pdnn_fc_ps_selftest(60, &l0, &l1);
            printf("selftest rc=%d loss %.4f -> %.4f %s\n", rc, l0, l1, rc ? pdnn_last_error() : "");
            return rc;
        }
        else { fprintf(stderr, "unknown argument %s (see --help)\n", a); return 2; }
    }
    pdnn_fc_ps_result_t r;
    int rc = pdnn_fc_ps_run(&c, &r);
    if (rc) { fprintf(stderr, "error: %s\n", pdnn_last_error()); return 1; }
    printf("%s: %d steps, %.3f ms/step, accepted=%lld stale=%lld abandoned=%lld avg_grads/step=%.2f loss %.4f -> %.4f err %.3f -> %.3f\n",
           r.scheme, r.steps, r.mean_step_ms, r.gradients_accepted, r.gradients_dropped_stale, r.steps_abandoned, r.avg_gradients_per_step,
           r.first_loss, r.final_loss, r.first_err, r.final_err);
    return 0;
}
