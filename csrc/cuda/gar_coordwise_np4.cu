#include "gar_coordwise.cuh"
namespace gb {
GB_DEFINE_COORDWISE_DISPATCH(4, 4, 4)
}
