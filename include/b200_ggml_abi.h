/* b200_ggml_abi.h -- the slice of the ggml ABI the B200 backend is called through.
 *
 * The drop-in boundary is the reference's ggml_backend interface: the host program hands the
 * backend `struct ggml_tensor` / `struct ggml_cgraph` objects it built itself and calls through five vtables.
 * This header restates the binary layout of exactly those objects so that libkcpp_b200.so builds with no
 * reference checkout present.  It is an ABI description, not a copy of the reference headers: only the structs
 * that cross the boundary, and only the enum VALUES the backend dispatches on.
 *
 * Every definition cites the reference declaration it mirrors (paths under /root/reference/ggml/), and
 * tests/cpp/abi_probe.cpp is compiled once against this file and once against the real headers and
 * tests/test_abi.py diffs every size, offset and enum value.  A maintainer integrating into a ggml tree compiles the backend with -DB200_USE_GGML_HEADERS instead,
 * which makes b200_ggml_backend.cpp include the tree's own ggml.h / ggml-backend-impl.h and skip this file.
 */
#ifndef B200_GGML_ABI_H
#define B200_GGML_ABI_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* include/ggml.h:218-227 */
#define GGML_MAX_DIMS      4
#define GGML_MAX_SRC       10
#define GGML_MAX_OP_PARAMS 64
#ifndef GGML_MAX_NAME
#define GGML_MAX_NAME      128 /* this fork: 128 (upstream 64) */
#endif
#define GGML_KQ_MASK_PAD   32  /* include/ggml.h:1802 */
#define GGML_ROPE_TYPE_NEOX 2  /* include/ggml.h:241 */

/* enum ggml_status, include/ggml.h:331-336 */
enum ggml_status {
    GGML_STATUS_ALLOC_FAILED = -2,
    GGML_STATUS_FAILED       = -1,
    GGML_STATUS_SUCCESS      = 0,
    GGML_STATUS_ABORTED      = 1,
};

/* enum ggml_type, include/ggml.h:361-398 -- only the values this backend stores or computes on */
enum ggml_type {
    GGML_TYPE_F32  = 0,
    GGML_TYPE_F16  = 1,
    GGML_TYPE_Q4_0 = 2,
    GGML_TYPE_Q4_1 = 3,
    GGML_TYPE_Q5_0 = 6,
    GGML_TYPE_Q5_1 = 7,
    GGML_TYPE_Q8_0 = 8,
    GGML_TYPE_Q8_1 = 9,
    GGML_TYPE_Q2_K = 10,
    GGML_TYPE_Q3_K = 11,
    GGML_TYPE_Q4_K = 12,
    GGML_TYPE_Q5_K = 13,
    GGML_TYPE_Q6_K = 14,
    GGML_TYPE_Q8_K = 15,
    GGML_TYPE_IQ4_NL = 20,
    GGML_TYPE_IQ4_XS = 23,
    GGML_TYPE_I8   = 24,
    GGML_TYPE_I16  = 25,
    GGML_TYPE_I32  = 26,
    GGML_TYPE_I64  = 27,
    GGML_TYPE_F64  = 28,
    GGML_TYPE_BF16 = 30,
    GGML_TYPE_COUNT = 36,
};

enum ggml_prec { GGML_PREC_DEFAULT = 0, GGML_PREC_F32 = 1 }; /* include/ggml.h:401-404 */

/* enum ggml_op, include/ggml.h:426-518 -- positional values of the ops in a llama / mixtral graph */
enum ggml_op {
    GGML_OP_NONE           = 0,
    GGML_OP_DUP            = 1,
    GGML_OP_ADD            = 2,
    GGML_OP_SUB            = 5,
    GGML_OP_MUL            = 6,
    GGML_OP_DIV            = 7,
    GGML_OP_SQR            = 8,
    GGML_OP_SUM_ROWS       = 14,
    GGML_OP_ARGMAX         = 16,
    GGML_OP_RMS_NORM       = 23,
    GGML_OP_MUL_MAT        = 26,
    GGML_OP_MUL_MAT_ID     = 27,
    GGML_OP_SCALE          = 29,
    GGML_OP_CPY            = 31,
    GGML_OP_CONT           = 32,
    GGML_OP_RESHAPE        = 33,
    GGML_OP_VIEW           = 34,
    GGML_OP_PERMUTE        = 35,
    GGML_OP_TRANSPOSE      = 36,
    GGML_OP_GET_ROWS       = 37,
    GGML_OP_SOFT_MAX       = 42,
    GGML_OP_ROPE           = 44,
    GGML_OP_CLAMP          = 46,
    GGML_OP_ARGSORT        = 58,
    GGML_OP_FLASH_ATTN_EXT = 60,
    GGML_OP_UNARY          = 69,
    GGML_OP_MAP_CUSTOM1    = 75, /* carries the tensor-parallel collectives (b200_comm.h); llama.cpp never emits it */
    GGML_OP_COUNT          = 81,
};

/* enum ggml_unary_op, include/ggml.h:520-537 */
enum ggml_unary_op {
    GGML_UNARY_OP_TANH       = 4,
    GGML_UNARY_OP_RELU       = 6,
    GGML_UNARY_OP_SIGMOID    = 7,
    GGML_UNARY_OP_GELU       = 8,
    GGML_UNARY_OP_GELU_QUICK = 9,
    GGML_UNARY_OP_SILU       = 10,
};

enum ggml_sort_order { GGML_SORT_ORDER_ASC = 0, GGML_SORT_ORDER_DESC = 1 }; /* include/ggml.h:1780-1783 */

enum ggml_tensor_flag { /* include/ggml.h:576-581 */
    GGML_TENSOR_FLAG_INPUT  = 1,
    GGML_TENSOR_FLAG_OUTPUT = 2,
    GGML_TENSOR_FLAG_PARAM  = 4,
    GGML_TENSOR_FLAG_LOSS   = 8,
};

struct ggml_backend_buffer;

/* struct ggml_tensor, include/ggml.h:584-619 */
struct ggml_tensor {
    enum ggml_type type;
    int32_t        backend; /* deprecated enum ggml_backend_type */
    struct ggml_backend_buffer *buffer;
    int64_t ne[GGML_MAX_DIMS];
    size_t  nb[GGML_MAX_DIMS];
    enum ggml_op op;
    int32_t op_params[GGML_MAX_OP_PARAMS / sizeof(int32_t)];
    int32_t flags;
    struct ggml_tensor *grad;
    struct ggml_tensor *src[GGML_MAX_SRC];
    struct ggml_tensor *view_src;
    size_t              view_offs;
    void *data;
    char  name[GGML_MAX_NAME];
    void *extra;
};

/* struct ggml_hash_set + struct ggml_cgraph, src/ggml-impl.h:80-84, 183-195 */
struct ggml_hash_set {
    size_t    size;
    uint32_t *used;
    struct ggml_tensor **keys;
};
struct ggml_cgraph {
    int size;
    int n_nodes;
    int n_leafs;
    struct ggml_tensor **nodes;
    struct ggml_tensor **grads;
    struct ggml_tensor **leafs;
    struct ggml_hash_set visited_hash_set;
    int32_t order; /* enum ggml_cgraph_eval_order */
};

/* handles, include/ggml-backend.h:10-16; include/ggml.h:694-695 */
typedef struct ggml_backend_buffer_type *ggml_backend_buffer_type_t;
typedef struct ggml_backend_buffer      *ggml_backend_buffer_t;
typedef struct ggml_backend_event       *ggml_backend_event_t;
typedef struct ggml_backend             *ggml_backend_t;
typedef void                            *ggml_backend_graph_plan_t;
typedef struct ggml_backend_reg         *ggml_backend_reg_t;
typedef struct ggml_backend_device      *ggml_backend_dev_t;
typedef uint8_t  ggml_guid[16];
typedef ggml_guid *ggml_guid_t;

enum ggml_backend_buffer_usage { /* include/ggml-backend.h:35-39 */
    GGML_BACKEND_BUFFER_USAGE_ANY     = 0,
    GGML_BACKEND_BUFFER_USAGE_WEIGHTS = 1,
    GGML_BACKEND_BUFFER_USAGE_COMPUTE = 2,
};
enum ggml_backend_dev_type { /* include/ggml-backend.h:124-130 */
    GGML_BACKEND_DEVICE_TYPE_CPU      = 0,
    GGML_BACKEND_DEVICE_TYPE_GPU      = 1,
    GGML_BACKEND_DEVICE_TYPE_CPU_FULL = 2,
    GGML_BACKEND_DEVICE_TYPE_GPU_FULL = 3,
};
struct ggml_backend_dev_caps { bool async, host_buffer, buffer_from_host_ptr, events; }; /* :133-142 */
struct ggml_backend_dev_props { /* :145-152 */
    const char *name, *description;
    size_t memory_free, memory_total;
    enum ggml_backend_dev_type type;
    struct ggml_backend_dev_caps caps;
};
/* functions obtainable through reg->iface.get_proc_address, include/ggml-backend.h:171-173 */
typedef ggml_backend_buffer_type_t (*ggml_backend_split_buffer_type_t)(const float *);

/* the five vtables, src/ggml-backend-impl.h:15-27, 39-57, 86-126, 146-188, 200-210 */
struct ggml_backend_buffer_type_i {
    const char *          (*get_name)(ggml_backend_buffer_type_t buft);
    ggml_backend_buffer_t (*alloc_buffer)(ggml_backend_buffer_type_t buft, size_t size);
    size_t                (*get_alignment)(ggml_backend_buffer_type_t buft);
    size_t                (*get_max_size)(ggml_backend_buffer_type_t buft);
    size_t                (*get_alloc_size)(ggml_backend_buffer_type_t buft, const struct ggml_tensor *tensor);
    bool                  (*is_host)(ggml_backend_buffer_type_t buft);
};
struct ggml_backend_buffer_type { /* :29-33 */
    struct ggml_backend_buffer_type_i iface;
    ggml_backend_dev_t device;
    void *context;
};
struct ggml_backend_buffer_i {
    const char *(*get_name)(ggml_backend_buffer_t buffer);
    void  (*free_buffer)(ggml_backend_buffer_t buffer);
    void *(*get_base)(ggml_backend_buffer_t buffer);
    void  (*init_tensor)(ggml_backend_buffer_t buffer, struct ggml_tensor *tensor);
    void  (*memset_tensor)(ggml_backend_buffer_t buffer, struct ggml_tensor *tensor, uint8_t value, size_t offset, size_t size);
    void  (*set_tensor)(ggml_backend_buffer_t buffer, struct ggml_tensor *tensor, const void *data, size_t offset, size_t size);
    void  (*get_tensor)(ggml_backend_buffer_t buffer, const struct ggml_tensor *tensor, void *data, size_t offset, size_t size);
    bool  (*cpy_tensor)(ggml_backend_buffer_t buffer, const struct ggml_tensor *src, struct ggml_tensor *dst);
    void  (*clear)(ggml_backend_buffer_t buffer, uint8_t value);
    void  (*reset)(ggml_backend_buffer_t buffer);
};
struct ggml_backend_buffer { /* :59-65 */
    struct ggml_backend_buffer_i iface;
    ggml_backend_buffer_type_t   buft;
    void  *context;
    size_t size;
    enum ggml_backend_buffer_usage usage;
};
struct ggml_backend_i {
    const char *(*get_name)(ggml_backend_t backend);
    void (*free)(ggml_backend_t backend);
    ggml_backend_buffer_type_t (*get_default_buffer_type)(ggml_backend_t backend);
    void (*set_tensor_async)(ggml_backend_t backend, struct ggml_tensor *tensor, const void *data, size_t offset, size_t size);
    void (*get_tensor_async)(ggml_backend_t backend, const struct ggml_tensor *tensor, void *data, size_t offset, size_t size);
    bool (*cpy_tensor_async)(ggml_backend_t backend_src, ggml_backend_t backend_dst, const struct ggml_tensor *src, struct ggml_tensor *dst);
    void (*synchronize)(ggml_backend_t backend);
    ggml_backend_graph_plan_t (*graph_plan_create)(ggml_backend_t backend, const struct ggml_cgraph *cgraph);
    void (*graph_plan_free)(ggml_backend_t backend, ggml_backend_graph_plan_t plan);
    void (*graph_plan_update)(ggml_backend_t backend, ggml_backend_graph_plan_t plan, const struct ggml_cgraph *cgraph);
    enum ggml_status (*graph_plan_compute)(ggml_backend_t backend, ggml_backend_graph_plan_t plan);
    enum ggml_status (*graph_compute)(ggml_backend_t backend, struct ggml_cgraph *cgraph);
    bool (*supports_op)(ggml_backend_t backend, const struct ggml_tensor *op);
    bool (*supports_buft)(ggml_backend_t backend, ggml_backend_buffer_type_t buft);
    bool (*offload_op)(ggml_backend_t backend, const struct ggml_tensor *op);
    void (*event_record)(ggml_backend_t backend, ggml_backend_event_t event);
    void (*event_wait)(ggml_backend_t backend, ggml_backend_event_t event);
};
struct ggml_backend { /* :128-133 */
    ggml_guid_t guid;
    struct ggml_backend_i iface;
    ggml_backend_dev_t device;
    void *context;
};
struct ggml_backend_event { /* :135-138 */
    struct ggml_backend_device *device;
    void *context;
};
struct ggml_backend_device_i {
    const char *(*get_name)(ggml_backend_dev_t dev);
    const char *(*get_description)(ggml_backend_dev_t dev);
    void (*get_memory)(ggml_backend_dev_t dev, size_t *free, size_t *total);
    enum ggml_backend_dev_type (*get_type)(ggml_backend_dev_t dev);
    void (*get_props)(ggml_backend_dev_t dev, struct ggml_backend_dev_props *props);
    ggml_backend_t (*init_backend)(ggml_backend_dev_t dev, const char *params);
    ggml_backend_buffer_type_t (*get_buffer_type)(ggml_backend_dev_t dev);
    ggml_backend_buffer_type_t (*get_host_buffer_type)(ggml_backend_dev_t dev);
    ggml_backend_buffer_t (*buffer_from_host_ptr)(ggml_backend_dev_t dev, void *ptr, size_t size, size_t max_tensor_size);
    bool (*supports_op)(ggml_backend_dev_t dev, const struct ggml_tensor *op);
    bool (*supports_buft)(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft);
    bool (*offload_op)(ggml_backend_dev_t dev, const struct ggml_tensor *op);
    ggml_backend_event_t (*event_new)(ggml_backend_dev_t dev);
    void (*event_free)(ggml_backend_dev_t dev, ggml_backend_event_t event);
    void (*event_synchronize)(ggml_backend_dev_t dev, ggml_backend_event_t event);
};
struct ggml_backend_device { /* :190-194 */
    struct ggml_backend_device_i iface;
    ggml_backend_reg_t reg;
    void *context;
};
struct ggml_backend_reg_i {
    const char *(*get_name)(ggml_backend_reg_t reg);
    size_t (*get_device_count)(ggml_backend_reg_t reg);
    ggml_backend_dev_t (*get_device)(ggml_backend_reg_t reg, size_t index);
    void *(*get_proc_address)(ggml_backend_reg_t reg, const char *name);
};
struct ggml_backend_reg { /* :212-216 */
    struct ggml_backend_reg_i iface;
    void *context;
};

#ifdef __cplusplus
}
#endif
#endif /* B200_GGML_ABI_H */
