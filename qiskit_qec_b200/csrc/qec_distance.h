// qec_distance.h -- internal interface of qec_distance.cu (the C ABI wrappers live in qec_capi.cu)
#pragma once
#include <stddef.h>
#include <stdint.h>

// error codes: the values of QEC_ERR_* in include/qec_b200.h
#define QD_ERR_LOGIC (-1)
#define QD_ERR_INVALID (-2)
#define QD_ERR_CUDA (-3)
#define QD_ERR_LIMIT (-4)

// matrices are row-major uint8 [rows][2 n] in the reference's [X part | Z part] order
int qd_minimum_distance(int device, int sms, int n, int n_stab, const uint8_t *stab, int n_gauge,
                        const uint8_t *gauge, int max_weight, int *distance, char *err, size_t errlen);
int qd_minimum_distance_by_tests(int device, int sms, int n, int n_stab, const uint8_t *stab, int n_xl,
                                 const uint8_t *xl, int n_zl, const uint8_t *zl, int max_weight,
                                 int *distance, char *err, size_t errlen);
