#include "gar_coordwise.cuh"
namespace gb {
GB_DEFINE_COORDWISE_DISPATCH(12, 4, 4)
}
