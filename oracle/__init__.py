"""CPU oracle for the GPE / Helmholtz PINN train step.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package (`gpe_b200`) may import this
package; only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference`
legs of `bench.py` use it, and only as the checker / the timed CPU baseline.

What is here
------------
* `pinn_oracle`      -- plain-PyTorch (CPU, any dtype) restatement of the reference's PINN
                        classes: tanh-MLP forward, nested `torch.autograd.grad` derivatives,
                        every loss family of SURVEY.md section 8(a).  Each function cites the
                        reference file:line it follows.
* `jet_oracle`       -- per-point value / gradient / Hessian of the MLP and the parameter
                        gradient of a weighted jet functional, by autograd in fp64.
* `train_oracle`     -- the paper's curriculum driver (`train_gpe_model`, R/harmonic_pinn_simulation.py:220-431) on
                        the oracle classes with torch's own scheduler, to check `gpe_b200.training`.
* `reference_loader` -- imports the *real* reference classes from /root/reference (exists only
                        in the build container) to pin the restatement and to generate the
                        golden vectors under `tests/golden/` (script: tests/golden/make_golden.py).

Parity status: PINNED.  The restatement is checked against (a) outputs of the reference's own
classes run in the build container on seeded inputs (tests/golden/*.npz) and (b) the shipped
trained weights `harmonic_mode_zero_plot_data.pkl` / `box_mode_zero_plot_data.pkl`, whose
`mu_table` it reproduces (tests/test_oracle_golden.py).
"""
