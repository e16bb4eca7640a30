#include "mmvq_launch.cuh"
namespace b200 { B200_MMVQ_DEFINE(T_IQ4_NL) }
