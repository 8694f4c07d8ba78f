// TEMPORARY: CNN entry points not built yet in this commit; they fail loudly.
#include "nc_internal.h"
#define STUB(ret, name, args) extern "C" ret name args { nc_set_error(nullptr, NC_EUNSUPPORTED, #name ": CNN path not built yet"); return (ret)NC_EUNSUPPORTED; }
STUB(int, nc_net_create, (nc_ctx*, const char*, int, int, int, int, uint64_t, nc_net**))
STUB(int, nc_net_destroy, (nc_net*))
STUB(int64_t, nc_net_param_count, (nc_net*))
STUB(int, nc_net_tensor_count, (nc_net*))
STUB(int, nc_net_tensor_info, (nc_net*, int, int64_t*, int64_t*, int*, int64_t*))
extern "C" float* nc_net_params(nc_net*) { return nullptr; }
extern "C" float* nc_net_grads(nc_net*) { return nullptr; }
STUB(int, nc_net_set_params, (nc_net*, const float*))
STUB(int, nc_net_get_params, (nc_net*, float*))
STUB(int, nc_net_set_math, (nc_net*, int))
STUB(int, nc_net_forward, (nc_net*, const float*, int, float*, void*))
STUB(int, nc_net_predict, (nc_net*, const float*, int, int32_t*, void*))
STUB(int, nc_net_backward, (nc_net*, const float*, const float*, int, int, float, float*, int32_t*, void*))
STUB(int, nc_net_apply, (nc_net*, int, float, void*))
STUB(int, nc_net_train_step, (nc_net*, const float*, const float*, int, int, int, float, float*, int32_t*, void*))
STUB(int64_t, nc_net_global_step, (nc_net*))
STUB(int, nc_net_state_slots, (nc_net*, int))
STUB(int, nc_net_get_state, (nc_net*, int, float*))
STUB(int, nc_net_set_state, (nc_net*, int, const float*, int64_t))
