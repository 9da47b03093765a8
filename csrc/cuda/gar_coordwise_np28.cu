#include "gar_coordwise.cuh"
namespace gb {
GB_DEFINE_COORDWISE_DISPATCH(28, 1, 2)
}
