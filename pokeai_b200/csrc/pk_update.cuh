// DQN minibatch update kernels (trainer.py:152-198) -- filled in below.
#pragma once
#include "pk_common.cuh"
