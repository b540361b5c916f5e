# Convenience targets; everything they run is plain python / bash and can be called directly.
PY ?= python

.PHONY: build reference test gputest bench bench-ref smoke roofline clean

build:            ## compile csrc/*.cu for sm_100a into neuroimagedisttraining_b200/lib/libnidt_b200.so (no GPU needed)
	$(PY) __graft_entry__.py

reference:        ## install the unmodified reference into baseline/_ref (git-ignored) for the parity tests and the baseline arm
	bash scripts/install_reference.sh

test:             ## CPU suite (~6 min): parity with the reference, algorithms, CLIs, multi-rank gloo runs, property tests
	$(PY) -m pytest tests -x -q -m "not gpu"

gputest:          ## kernels and engines against PyTorch fp32 on a B200
	$(PY) -m pytest tests -x -q -m gpu

bench:            ## headline benchmark on one GPU (GPUS=N: torchrun over N GPUs)
ifeq ($(GPUS),)
	$(PY) bench.py --steps 5 --warmup 3
else
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(GPUS) --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $(GPUS) --steps 5 --warmup 3
endif

bench-ref:        ## the same benchmark through the unmodified reference
	$(PY) bench.py --impl reference --steps 3 --warmup 3

smoke:            ## two tiny rounds of all nine algorithms (GPUS=N for the client-parallel forms)
	bash scripts/jobs/smoke_all_algorithms.sh

roofline:         ## tables derived from the committed profiles (no GPU needed)
	$(PY) scripts/roofline.py profiles/profile_step_v18.txt --out profiles/roofline_step_v18.md
	$(PY) scripts/agg_model.py --out profiles/agg_model.md

clean:
	rm -rf neuroimagedisttraining_b200/build neuroimagedisttraining_b200/lib LOG results
