// ggml-b200-abi.h -- the slice of the ggml plugin ABI that ggml-b200 binds to.
//
// The backend is a drop-in for the reference's backend plugin interface
// (reference: ggml/src/ggml-backend-impl.h:17-210, ggml/include/ggml.h:389-431,469-576,655-687,
//  ggml/include/ggml-backend.h:49-53,130-171,200-212, ggml/src/ggml-impl.h:323-337).
// When the backend is built inside the reference tree (ggml/src/ggml-b200/)
// define GGML_B200_USE_GGML_HEADERS and the real headers are used.  Stand-alone (this repo,
// and the GPU box where /root/reference does not exist) the same binary layout is restated
// here: only what the backend reads or fills, in the reference's field order.  Every value
// and offset below is pinned against the reference headers by tests/test_abi.py.
#pragma once

#ifdef GGML_B200_USE_GGML_HEADERS
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#else

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGML_MAX_DIMS      4
#define GGML_MAX_SRC       10
#define GGML_MAX_OP_PARAMS 64
#define GGML_MAX_NAME      64
#define GGML_BACKEND_API_VERSION 2   // ggml-backend-impl.h:11

// ggml.h:389-431 -- only the tensor types the hot path touches
enum ggml_type {
    GGML_TYPE_F32  = 0,  GGML_TYPE_F16  = 1,  GGML_TYPE_Q4_0 = 2,  GGML_TYPE_Q4_1 = 3,
    GGML_TYPE_Q5_0 = 6,  GGML_TYPE_Q5_1 = 7,  GGML_TYPE_Q8_0 = 8,  GGML_TYPE_Q8_1 = 9,
    GGML_TYPE_Q2_K = 10, GGML_TYPE_Q3_K = 11, GGML_TYPE_Q4_K = 12, GGML_TYPE_Q5_K = 13,
    GGML_TYPE_Q6_K = 14, GGML_TYPE_Q8_K = 15, GGML_TYPE_I8 = 24,   GGML_TYPE_I16 = 25,
    GGML_TYPE_I32  = 26, GGML_TYPE_I64  = 27, GGML_TYPE_F64 = 28,  GGML_TYPE_BF16 = 30,
    GGML_TYPE_COUNT = 40,
};

enum ggml_prec { GGML_PREC_DEFAULT = 0, GGML_PREC_F32 = 10 };     // ggml.h:434-437

// ggml.h:469-576 -- numeric values of the ops this backend accepts (others are declined)
enum ggml_op {
    GGML_OP_NONE = 0, GGML_OP_DUP = 1, GGML_OP_ADD = 2, GGML_OP_ADD_ID = 3, GGML_OP_SUB = 6,
    GGML_OP_MUL = 7, GGML_OP_DIV = 8, GGML_OP_SUM_ROWS = 15, GGML_OP_RMS_NORM = 25,
    GGML_OP_MUL_MAT = 29, GGML_OP_MUL_MAT_ID = 30, GGML_OP_SCALE = 32, GGML_OP_CPY = 34,
    GGML_OP_CONT = 35, GGML_OP_RESHAPE = 36, GGML_OP_VIEW = 37, GGML_OP_PERMUTE = 38,
    GGML_OP_TRANSPOSE = 39, GGML_OP_GET_ROWS = 40, GGML_OP_SET_ROWS = 42, GGML_OP_SOFT_MAX = 46,
    GGML_OP_ROPE = 48, GGML_OP_CLAMP = 50, GGML_OP_ARGSORT = 68, GGML_OP_TOP_K = 69,
    GGML_OP_FLASH_ATTN_EXT = 73, GGML_OP_UNARY = 85, GGML_OP_GLU = 94, GGML_OP_COUNT = 95,
};

enum ggml_unary_op { GGML_UNARY_OP_SILU = 10 };                    // ggml.h:578-603
enum ggml_glu_op   { GGML_GLU_OP_REGLU = 0, GGML_GLU_OP_GEGLU = 1, GGML_GLU_OP_SWIGLU = 2 };  // ggml.h:605-614

enum ggml_status {                                                 // ggml.h:358-363
    GGML_STATUS_ALLOC_FAILED = -2, GGML_STATUS_FAILED = -1, GGML_STATUS_SUCCESS = 0, GGML_STATUS_ABORTED = 1,
};

typedef uint16_t ggml_fp16_t;
typedef uint8_t  ggml_guid[16];
typedef ggml_guid * ggml_guid_t;

struct ggml_backend_buffer;

// ggml.h:655-687 (336 bytes)
struct ggml_tensor {
    enum ggml_type type;
    struct ggml_backend_buffer * buffer;
    int64_t ne[GGML_MAX_DIMS];
    size_t  nb[GGML_MAX_DIMS];
    enum ggml_op op;
    int32_t op_params[GGML_MAX_OP_PARAMS / sizeof(int32_t)];
    int32_t flags;
    struct ggml_tensor * src[GGML_MAX_SRC];
    struct ggml_tensor * view_src;
    size_t               view_offs;
    void * data;
    char name[GGML_MAX_NAME];
    void * extra;
    char padding[8];
};

// ggml-impl.h:323-337 -- the backend only walks nodes[0..n_nodes)
struct ggml_cgraph {
    int size;
    int n_nodes;
    int n_leafs;
    struct ggml_tensor ** nodes;
    struct ggml_tensor ** grads;
    struct ggml_tensor ** grad_accs;
    struct ggml_tensor ** leafs;
    int32_t * use_counts;
    /* struct ggml_hash_set visited_hash_set; enum ggml_cgraph_eval_order order;  -- not read */
};

typedef struct ggml_backend_buffer_type * ggml_backend_buffer_type_t;
typedef struct ggml_backend_buffer      * ggml_backend_buffer_t;
typedef struct ggml_backend_event       * ggml_backend_event_t;
typedef struct ggml_backend             * ggml_backend_t;
typedef void                            * ggml_backend_graph_plan_t;
typedef struct ggml_backend_reg         * ggml_backend_reg_t;
typedef struct ggml_backend_device      * ggml_backend_dev_t;

enum ggml_backend_buffer_usage {                                   // ggml-backend.h:49-53
    GGML_BACKEND_BUFFER_USAGE_ANY = 0, GGML_BACKEND_BUFFER_USAGE_WEIGHTS = 1, GGML_BACKEND_BUFFER_USAGE_COMPUTE = 2,
};
enum ggml_backend_dev_type {                                       // ggml-backend.h:130-139
    GGML_BACKEND_DEVICE_TYPE_CPU, GGML_BACKEND_DEVICE_TYPE_GPU, GGML_BACKEND_DEVICE_TYPE_IGPU, GGML_BACKEND_DEVICE_TYPE_ACCEL,
};
struct ggml_backend_dev_caps { bool async; bool host_buffer; bool buffer_from_host_ptr; bool events; };
struct ggml_backend_dev_props {                                    // ggml-backend.h:154-171
    const char * name; const char * description; size_t memory_free; size_t memory_total;
    enum ggml_backend_dev_type type; const char * device_id; struct ggml_backend_dev_caps caps;
};
struct ggml_backend_feature { const char * name; const char * value; };   // ggml-backend.h:208-211

// ---- vtables, ggml-backend-impl.h ------------------------------------------------------
struct ggml_backend_buffer_type_i {                                // :17-29
    const char *          (*get_name)      (ggml_backend_buffer_type_t buft);
    ggml_backend_buffer_t (*alloc_buffer)  (ggml_backend_buffer_type_t buft, size_t size);
    size_t                (*get_alignment) (ggml_backend_buffer_type_t buft);
    size_t                (*get_max_size)  (ggml_backend_buffer_type_t buft);
    size_t                (*get_alloc_size)(ggml_backend_buffer_type_t buft, const struct ggml_tensor * tensor);
    bool                  (*is_host)       (ggml_backend_buffer_type_t buft);
};
struct ggml_backend_buffer_type { struct ggml_backend_buffer_type_i iface; ggml_backend_dev_t device; void * context; };

struct ggml_backend_buffer_i {                                     // :41-58
    void             (*free_buffer)  (ggml_backend_buffer_t buffer);
    void *           (*get_base)     (ggml_backend_buffer_t buffer);
    enum ggml_status (*init_tensor)  (ggml_backend_buffer_t buffer, struct ggml_tensor * tensor);
    void             (*memset_tensor)(ggml_backend_buffer_t buffer, struct ggml_tensor * tensor, uint8_t value, size_t offset, size_t size);
    void             (*set_tensor)   (ggml_backend_buffer_t buffer, struct ggml_tensor * tensor, const void * data, size_t offset, size_t size);
    void             (*get_tensor)   (ggml_backend_buffer_t buffer, const struct ggml_tensor * tensor, void * data, size_t offset, size_t size);
    bool             (*cpy_tensor)   (ggml_backend_buffer_t buffer, const struct ggml_tensor * src, struct ggml_tensor * dst);
    void             (*clear)        (ggml_backend_buffer_t buffer, uint8_t value);
    void             (*reset)        (ggml_backend_buffer_t buffer);
};
struct ggml_backend_buffer {                                       // :60-66
    struct ggml_backend_buffer_i iface; ggml_backend_buffer_type_t buft; void * context; size_t size;
    enum ggml_backend_buffer_usage usage;
};

struct ggml_backend_i {                                            // :87-120
    const char * (*get_name)(ggml_backend_t backend);
    void (*free)(ggml_backend_t backend);
    void (*set_tensor_async)(ggml_backend_t backend, struct ggml_tensor * tensor, const void * data, size_t offset, size_t size);
    void (*get_tensor_async)(ggml_backend_t backend, const struct ggml_tensor * tensor, void * data, size_t offset, size_t size);
    bool (*cpy_tensor_async)(ggml_backend_t backend_src, ggml_backend_t backend_dst, const struct ggml_tensor * src, struct ggml_tensor * dst);
    void (*synchronize)(ggml_backend_t backend);
    ggml_backend_graph_plan_t (*graph_plan_create) (ggml_backend_t backend, const struct ggml_cgraph * cgraph);
    void                      (*graph_plan_free)   (ggml_backend_t backend, ggml_backend_graph_plan_t plan);
    void                      (*graph_plan_update) (ggml_backend_t backend, ggml_backend_graph_plan_t plan, const struct ggml_cgraph * cgraph);
    enum ggml_status          (*graph_plan_compute)(ggml_backend_t backend, ggml_backend_graph_plan_t plan);
    enum ggml_status          (*graph_compute)     (ggml_backend_t backend, struct ggml_cgraph * cgraph);
    void (*event_record)(ggml_backend_t backend, ggml_backend_event_t event);
    void (*event_wait)  (ggml_backend_t backend, ggml_backend_event_t event);
    void (*graph_optimize)(ggml_backend_t backend, struct ggml_cgraph * cgraph);
};
struct ggml_backend { ggml_guid_t guid; struct ggml_backend_i iface; ggml_backend_dev_t device; void * context; };   // :122-127
struct ggml_backend_event { struct ggml_backend_device * device; void * context; };                                  // :129-132

struct ggml_backend_device_i {                                     // :140-182
    const char * (*get_name)(ggml_backend_dev_t dev);
    const char * (*get_description)(ggml_backend_dev_t dev);
    void         (*get_memory)(ggml_backend_dev_t dev, size_t * free, size_t * total);
    enum ggml_backend_dev_type (*get_type)(ggml_backend_dev_t dev);
    void (*get_props)(ggml_backend_dev_t dev, struct ggml_backend_dev_props * props);
    ggml_backend_t (*init_backend)(ggml_backend_dev_t dev, const char * params);
    ggml_backend_buffer_type_t (*get_buffer_type)(ggml_backend_dev_t dev);
    ggml_backend_buffer_type_t (*get_host_buffer_type)(ggml_backend_dev_t dev);
    ggml_backend_buffer_t (*buffer_from_host_ptr)(ggml_backend_dev_t dev, void * ptr, size_t size, size_t max_tensor_size);
    bool (*supports_op)(ggml_backend_dev_t dev, const struct ggml_tensor * op);
    bool (*supports_buft)(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft);
    bool (*offload_op)(ggml_backend_dev_t dev, const struct ggml_tensor * op);
    ggml_backend_event_t (*event_new)         (ggml_backend_dev_t dev);
    void                 (*event_free)        (ggml_backend_dev_t dev, ggml_backend_event_t event);
    void                 (*event_synchronize) (ggml_backend_dev_t dev, ggml_backend_event_t event);
};
struct ggml_backend_device { struct ggml_backend_device_i iface; ggml_backend_reg_t reg; void * context; };          // :184-188

struct ggml_backend_reg_i {                                        // :194-204
    const char *       (*get_name)(ggml_backend_reg_t reg);
    size_t             (*get_device_count)(ggml_backend_reg_t reg);
    ggml_backend_dev_t (*get_device)(ggml_backend_reg_t reg, size_t index);
    void *             (*get_proc_address)(ggml_backend_reg_t reg, const char * name);
};
struct ggml_backend_reg { int api_version; struct ggml_backend_reg_i iface; void * context; };                       // :206-210

// proc-address signatures callers ask for by name (ggml-backend.h:200-212)
typedef ggml_backend_buffer_type_t   (*ggml_backend_split_buffer_type_t)(int main_device, const float * tensor_split);
typedef ggml_backend_buffer_type_t * (*ggml_backend_dev_get_extra_bufts_t)(ggml_backend_dev_t device);
typedef struct ggml_backend_feature * (*ggml_backend_get_features_t)(ggml_backend_reg_t reg);

#ifdef __cplusplus
}
#endif
#endif // GGML_B200_USE_GGML_HEADERS
