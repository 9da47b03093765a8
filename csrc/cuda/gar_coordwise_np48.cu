#include "gar_coordwise.cuh"
namespace gb {
GB_DEFINE_COORDWISE_DISPATCH(48, 1, 1)
}
