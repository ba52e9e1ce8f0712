"""CPU oracle for the data-parallel training step of the reference's ``imagenet.py``.

THIS PACKAGE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker.  Nothing under
``imagent_distributed_training_pytorch_with_slurm_b200/`` imports it; the product
path raises when its CUDA library is missing instead of falling back to this code.

What it restates (numpy / plain torch-CPU, one function per reference site):

==================  ==========================================================
module              reference site it follows
==================  ==========================================================
``step_helpers``    ``imagenet.py:44-60`` AverageMeter, ``:63-79`` accuracy,
                    ``:82-87`` reduce_tensor, ``:154-162`` adjust_learning_rate
``sampler``         ``imagenet.py:346-347,375`` -> torch DistributedSampler
                    (``torch/utils/data/distributed.py:107-141``)
``input_pipeline``  ``imagenet.py:280-283`` Resize/ToTensor/Normalize (PIL's
                    fixed-point bilinear resample, Pillow 12.2 ``Resample.c``)
                    and torchvision ``RandomResizedCrop.get_params`` /
                    ``RandomHorizontalFlip`` (``transforms.py:928-981,712-722``)
``sgd``             ``imagenet.py:325,131`` -> ``torch/optim/sgd.py:343-380``
``batchnorm``       ``torchvision/models/resnet.py:93-103,147-161,240-243`` ->
                    ``torch/nn/modules/batchnorm.py:160-206``
``maxpool``         ``torchvision/models/resnet.py:200,270`` MaxPool2d(3,2,1) and its
                    gradient (ATen first-maximum tie rule)
``cross_entropy``   ``imagenet.py:124,323`` -> ``torch/nn/functional.py:3421-3504``
                    fused with ``accuracy`` (``imagenet.py:63-79``)
``train_cpu``       ``imagenet.py:97-151`` train() and ``:166-210`` validate()
                    with only the four device calls removed, stock torch
                    DDP/ResNet/CrossEntropy/SGD over gloo
==================  ==========================================================

Parity status: the reference ships no tests, golden vectors or fixtures
(SURVEY.md section 4), and the arithmetic lives in un-vendored, un-pinned third
party code (torch / torchvision / Pillow).  The oracle is therefore pinned
against (a) the reference's own helper functions imported in the build container
through ``oracle/ref_shim.py`` and (b) the installed torch 2.11.0 / torchvision
0.26.0 / Pillow 12.2.0 run on the same seeded inputs; the resulting vectors are
committed under ``tests/golden/`` together with ``oracle/make_golden.py`` that
produced them.  The LR-schedule values are additionally pinned by the
reference's published log (``imagent_sgd.out:274,454,634,814``).
"""
