/* JNI shim over include/photon_ml_b200.h: one native per method of com.linkedin.photon.ml.function.glm.NativeGLM
 * (jni/NativeGLM.scala), each a thin wrapper around one C-ABI call.
 *
 * No JVM ships in the build image, so this file is not linked there; tests/test_jni_text.py compiles it
 * (gcc -fsyntax-only -Wall -Werror) against tests/jni_stub/jni.h, a stand-in that declares the JNI types and the
 * JNIEnv functions used below, and checks that every @native of NativeGLM.scala has its function here -- signature
 * drift against include/photon_ml_b200.h fails the CPU test suite.  Build for real with
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude jni/photon_ml_jni.c \
 *       -Lphoton_ml_b200/lib -lphoton_ml_b200 -o libphoton_ml_b200_jni.so
 *
 * Array access: calls that PACK data (datasetCreate, batchCreate, set*) allocate device memory, sort and
 * synchronise, so they take copies of the Java arrays with Get<Type>ArrayElements -- never inside a
 * GetPrimitiveArrayCritical region, where the JVM may be holding off garbage collection for every thread.
 * Evaluation calls move two d-vectors and allocate nothing once warmUp() has run on the data set, so they pin the
 * arrays with GetPrimitiveArrayCritical and avoid the copies.
 * Errors: PML_EINVAL -> IllegalArgumentException (the reference's `require`), anything else -> RuntimeException. */
#include <jni.h>
#include <stdint.h>
#include <stddef.h>

#include "photon_ml_b200.h"

#define NATIVE(ret, name) JNIEXPORT ret JNICALL Java_com_linkedin_photon_ml_function_glm_NativeGLM_##name

static void throw_for(JNIEnv *env, int rc)
{
    const char *cls = rc == PML_EINVAL ? "java/lang/IllegalArgumentException" : "java/lang/RuntimeException";
    jclass c = (*env)->FindClass(env, cls);
    if (c != NULL) (*env)->ThrowNew(env, c, pml_last_error());
}

static void throw_msg(JNIEnv *env, const char *cls, const char *msg)
{
    jclass c = (*env)->FindClass(env, cls);
    if (c != NULL) (*env)->ThrowNew(env, c, msg);
}

#define CTX(h) ((pml_ctx *)(intptr_t)(h))
#define DS(h) ((pml_dataset *)(intptr_t)(h))
#define BATCH(h) ((pml_batch *)(intptr_t)(h))
#define PEER(h) ((pml_peer *)(intptr_t)(h))

/* copies of Java arrays for the packing calls; a NULL Java array stays NULL (offsets / weights are optional) */
static jlong *longs(JNIEnv *env, jlongArray a) { return a ? (*env)->GetLongArrayElements(env, a, NULL) : NULL; }
static jint *ints(JNIEnv *env, jintArray a) { return a ? (*env)->GetIntArrayElements(env, a, NULL) : NULL; }
static jdouble *doubles(JNIEnv *env, jdoubleArray a) { return a ? (*env)->GetDoubleArrayElements(env, a, NULL) : NULL; }
static jbyte *bytes(JNIEnv *env, jbyteArray a) { return a ? (*env)->GetByteArrayElements(env, a, NULL) : NULL; }
static void free_longs(JNIEnv *env, jlongArray a, jlong *p) { if (p) (*env)->ReleaseLongArrayElements(env, a, p, JNI_ABORT); }
static void free_ints(JNIEnv *env, jintArray a, jint *p) { if (p) (*env)->ReleaseIntArrayElements(env, a, p, JNI_ABORT); }
static void free_doubles(JNIEnv *env, jdoubleArray a, jdouble *p) { if (p) (*env)->ReleaseDoubleArrayElements(env, a, p, JNI_ABORT); }
static void free_bytes(JNIEnv *env, jbyteArray a, jbyte *p) { if (p) (*env)->ReleaseByteArrayElements(env, a, p, JNI_ABORT); }

/* pinned views for the evaluation calls */
static double *pin(JNIEnv *env, jdoubleArray a) { return a ? (double *)(*env)->GetPrimitiveArrayCritical(env, a, NULL) : NULL; }
static void unpin_in(JNIEnv *env, jdoubleArray a, double *p) { if (p) (*env)->ReleasePrimitiveArrayCritical(env, a, p, JNI_ABORT); }
static void unpin_out(JNIEnv *env, jdoubleArray a, double *p) { if (p) (*env)->ReleasePrimitiveArrayCritical(env, a, p, 0); }

/* ---- contexts ------------------------------------------------------------------------------------------------ */
NATIVE(jlong, ctxCreate)(JNIEnv *env, jclass cls, jint device)
{
    (void)cls;
    pml_ctx *ctx = NULL;
    int rc = pml_ctx_create((int)device, &ctx);
    if (rc) { throw_for(env, rc); return 0; }
    return (jlong)(intptr_t)ctx;
}

NATIVE(void, ctxDestroy)(JNIEnv *env, jclass cls, jlong ctx)
{
    (void)cls;
    int rc = pml_ctx_destroy(CTX(ctx));
    if (rc) throw_for(env, rc);
}

/* ---- data sets: RDD[LabeledPoint] partitions / Iterable[LabeledPoint] flattened to CSR on the Scala side -------- */
NATIVE(jlong, datasetCreate)(JNIEnv *env, jclass cls, jlong ctx, jlong n, jint d, jlongArray rowPtr, jintArray col,
                             jdoubleArray val, jdoubleArray labels, jdoubleArray offsets, jdoubleArray weights)
{
    (void)cls;
    if (rowPtr == NULL || labels == NULL || (col == NULL) != (val == NULL)) {
        throw_msg(env, "java/lang/IllegalArgumentException", "rowPtr, labels and (col, value) are required");
        return 0;
    }
    jlong *rp = longs(env, rowPtr);
    jint *ci = ints(env, col);
    jdouble *v = doubles(env, val), *y = doubles(env, labels), *o = doubles(env, offsets), *w = doubles(env, weights);
    pml_dataset *ds = NULL;
    int rc = pml_dataset_create_csr(CTX(ctx), (int64_t)n, (int32_t)d, (const int64_t *)rp, (const int32_t *)ci, v, PML_F64,
                                    PML_STORE_AUTO, y, o, w, &ds);
    free_doubles(env, weights, w);
    free_doubles(env, offsets, o);
    free_doubles(env, labels, y);
    free_doubles(env, val, v);
    free_ints(env, col, ci);
    free_longs(env, rowPtr, rp);
    if (rc) { throw_for(env, rc); return 0; }
    return (jlong)(intptr_t)ds;
}

NATIVE(void, datasetDestroy)(JNIEnv *env, jclass cls, jlong ds)
{
    (void)cls;
    int rc = pml_dataset_destroy(DS(ds));
    if (rc) throw_for(env, rc);
}

NATIVE(void, setOffsets)(JNIEnv *env, jclass cls, jlong ds, jdoubleArray offsets)
{
    (void)cls;
    jdouble *o = doubles(env, offsets);   /* NULL clears the offsets */
    int rc = pml_dataset_set_offsets(DS(ds), o);
    free_doubles(env, offsets, o);
    if (rc) throw_for(env, rc);
}

NATIVE(void, setWeights)(JNIEnv *env, jclass cls, jlong ds, jdoubleArray weights)
{
    (void)cls;
    jdouble *w = doubles(env, weights);
    int rc = pml_dataset_set_weights(DS(ds), w);
    free_doubles(env, weights, w);
    if (rc) throw_for(env, rc);
}

NATIVE(void, setNormalization)(JNIEnv *env, jclass cls, jlong ds, jdoubleArray factors, jdoubleArray shifts, jint interceptId)
{
    (void)cls;
    jdouble *f = doubles(env, factors), *s = doubles(env, shifts);
    int rc = pml_dataset_set_normalization(DS(ds), f, s, (int32_t)interceptId);
    free_doubles(env, shifts, s);
    free_doubles(env, factors, f);
    if (rc) throw_for(env, rc);
}

/* One evaluation of every kind, so that the lazily allocated buffers (Hessian-vector cache, score staging) exist
 * before the first call that pins Java arrays. */
NATIVE(void, warmUp)(JNIEnv *env, jclass cls, jlong ds, jint loss, jint d, jlong n)
{
    (void)cls;
    jdoubleArray zero = (*env)->NewDoubleArray(env, (jsize)d), out = (*env)->NewDoubleArray(env, (jsize)d);
    jdoubleArray rows = (*env)->NewDoubleArray(env, (jsize)n);
    if (zero == NULL || out == NULL || rows == NULL) return;   /* OutOfMemoryError is pending */
    jdouble *z = doubles(env, zero), *g = doubles(env, out), *r = doubles(env, rows);
    double value = 0.0;
    int rc = pml_value_and_gradient(DS(ds), (int)loss, z, &value, g);
    if (rc == PML_OK && loss != PML_LOSS_SMOOTHED_HINGE) rc = pml_hessian_vector(DS(ds), (int)loss, z, 1, z, g);
    if (rc == PML_OK) rc = pml_scores(DS(ds), z, r);
    free_doubles(env, rows, r);
    free_doubles(env, out, g);
    free_doubles(env, zero, z);
    if (rc) throw_for(env, rc);
}

/* ---- DistributedGLMLossFunction / SingleNodeGLMLossFunction bodies --------------------------------------------- */
NATIVE(jdouble, valueAndGradient)(JNIEnv *env, jclass cls, jlong ds, jint loss, jdoubleArray coef, jdoubleArray outGrad)
{
    (void)cls;
    double value = 0.0;
    double *w = pin(env, coef), *g = pin(env, outGrad);   /* outGrad == null: the value only */
    int rc = w ? pml_value_and_gradient(DS(ds), (int)loss, w, &value, g) : PML_EINVAL;
    unpin_out(env, outGrad, g);
    unpin_in(env, coef, w);
    if (!w) { throw_msg(env, "java/lang/IllegalArgumentException", "null coefficients"); return 0.0; }
    if (rc) throw_for(env, rc);
    return (jdouble)value;
}

NATIVE(void, hessianVector)(JNIEnv *env, jclass cls, jlong ds, jint loss, jdoubleArray coef, jlong tag, jdoubleArray v,
                            jdoubleArray out)
{
    (void)cls;
    double *w = pin(env, coef), *x = pin(env, v), *h = pin(env, out);
    int rc = (x && h) ? pml_hessian_vector(DS(ds), (int)loss, w, (uint64_t)tag, x, h) : PML_EINVAL;
    unpin_out(env, out, h);
    unpin_in(env, v, x);
    unpin_in(env, coef, w);
    if (!(x && h)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

NATIVE(void, hessianDiagonal)(JNIEnv *env, jclass cls, jlong ds, jint loss, jdoubleArray coef, jdoubleArray out)
{
    (void)cls;
    double *w = pin(env, coef), *h = pin(env, out);
    int rc = (w && h) ? pml_hessian_diagonal(DS(ds), (int)loss, w, h) : PML_EINVAL;
    unpin_out(env, out, h);
    unpin_in(env, coef, w);
    if (!(w && h)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

NATIVE(void, scores)(JNIEnv *env, jclass cls, jlong ds, jdoubleArray coef, jdoubleArray outScores)
{
    (void)cls;
    double *w = pin(env, coef), *z = pin(env, outScores);
    int rc = (w && z) ? pml_scores(DS(ds), w, z) : PML_EINVAL;
    unpin_out(env, outScores, z);
    unpin_in(env, coef, w);
    if (!(w && z)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

/* ---- random-effect coordinates: all entities of a partition in one call ------------------------------------------ */
NATIVE(jlong, batchCreate)(JNIEnv *env, jclass cls, jlong ctx, jlongArray entRowPtr, jintArray entDim, jlongArray rowPtr,
                           jintArray colLocal, jdoubleArray val, jdoubleArray labels, jdoubleArray offsets,
                           jdoubleArray weights)
{
    (void)cls;
    if (entRowPtr == NULL || entDim == NULL || rowPtr == NULL || labels == NULL) {
        throw_msg(env, "java/lang/IllegalArgumentException", "entRowPtr, entDim, rowPtr and labels are required");
        return 0;
    }
    const jsize n_ent = (*env)->GetArrayLength(env, entDim);
    jlong *erp = longs(env, entRowPtr), *rp = longs(env, rowPtr);
    jint *ed = ints(env, entDim), *ci = ints(env, colLocal);
    jdouble *v = doubles(env, val), *y = doubles(env, labels), *o = doubles(env, offsets), *w = doubles(env, weights);
    pml_batch *b = NULL;
    int rc = pml_batch_create_csr(CTX(ctx), (int64_t)n_ent, (const int64_t *)erp, (const int32_t *)ed, (const int64_t *)rp,
                                  (const int32_t *)ci, v, PML_F64, PML_STORE_AUTO, y, o, w, &b);
    free_doubles(env, weights, w);
    free_doubles(env, offsets, o);
    free_doubles(env, labels, y);
    free_doubles(env, val, v);
    free_ints(env, colLocal, ci);
    free_ints(env, entDim, ed);
    free_longs(env, rowPtr, rp);
    free_longs(env, entRowPtr, erp);
    if (rc) { throw_for(env, rc); return 0; }
    return (jlong)(intptr_t)b;
}

NATIVE(void, batchDestroy)(JNIEnv *env, jclass cls, jlong batch)
{
    (void)cls;
    int rc = pml_batch_destroy(BATCH(batch));
    if (rc) throw_for(env, rc);
}

NATIVE(void, batchValueAndGradient)(JNIEnv *env, jclass cls, jlong batch, jint loss, jdoubleArray coefs,
                                    jdoubleArray outValues, jdoubleArray outGrads)
{
    (void)cls;
    double *w = pin(env, coefs), *f = pin(env, outValues), *g = pin(env, outGrads);
    int rc = (w && f) ? pml_batch_value_and_gradient(BATCH(batch), (int)loss, w, f, g) : PML_EINVAL;
    unpin_out(env, outGrads, g);
    unpin_out(env, outValues, f);
    unpin_in(env, coefs, w);
    if (!(w && f)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

NATIVE(void, batchHessianVector)(JNIEnv *env, jclass cls, jlong batch, jint loss, jdoubleArray coefs, jlong tag,
                                 jdoubleArray vs, jdoubleArray outHvs)
{
    (void)cls;
    double *w = pin(env, coefs), *x = pin(env, vs), *h = pin(env, outHvs);
    int rc = (x && h) ? pml_batch_hessian_vector(BATCH(batch), (int)loss, w, (uint64_t)tag, x, h) : PML_EINVAL;
    unpin_out(env, outHvs, h);
    unpin_in(env, vs, x);
    unpin_in(env, coefs, w);
    if (!(x && h)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

NATIVE(void, batchSetOffsets)(JNIEnv *env, jclass cls, jlong batch, jdoubleArray offsets)
{
    (void)cls;
    jdouble *o = doubles(env, offsets);
    int rc = pml_batch_set_offsets(BATCH(batch), o);
    free_doubles(env, offsets, o);
    if (rc) throw_for(env, rc);
}

NATIVE(void, batchSetActive)(JNIEnv *env, jclass cls, jlong batch, jbyteArray active)
{
    (void)cls;
    jbyte *a = bytes(env, active);   /* NULL: every entity active again */
    int rc = pml_batch_set_active(BATCH(batch), (const uint8_t *)a);
    free_bytes(env, active, a);
    if (rc) throw_for(env, rc);
}

/* ---- several executors, one GPU each: the exchange that replaces broadcast + treeAggregate ------------------------ */
NATIVE(jlong, peerCreate)(JNIEnv *env, jclass cls, jlong ctx, jint rank, jint world, jint d)
{
    (void)cls;
    pml_peer *p = NULL;
    int rc = pml_peer_create(CTX(ctx), (int)rank, (int)world, (int32_t)d, &p);
    if (rc) { throw_for(env, rc); return 0; }
    return (jlong)(intptr_t)p;
}

NATIVE(jbyteArray, peerHandle)(JNIEnv *env, jclass cls, jlong peer)
{
    (void)cls;
    jbyte handle[PML_PEER_HANDLE_BYTES];
    int rc = pml_peer_handle(PEER(peer), handle);
    if (rc) { throw_for(env, rc); return NULL; }
    jbyteArray out = (*env)->NewByteArray(env, PML_PEER_HANDLE_BYTES);
    if (out != NULL) (*env)->SetByteArrayRegion(env, out, 0, PML_PEER_HANDLE_BYTES, handle);
    return out;
}

NATIVE(void, peerConnect)(JNIEnv *env, jclass cls, jlong peer, jbyteArray allHandles)
{
    (void)cls;
    jbyte *h = bytes(env, allHandles);
    int rc = h ? pml_peer_connect(PEER(peer), h) : PML_EINVAL;
    free_bytes(env, allHandles, h);
    if (!h) { throw_msg(env, "java/lang/IllegalArgumentException", "null handles"); return; }
    if (rc) throw_for(env, rc);
}

/* [lo, hi) of the coefficient vector this rank uploads and of the result it downloads: out = long[2] */
NATIVE(void, peerSlice)(JNIEnv *env, jclass cls, jlong peer, jlongArray out)
{
    (void)cls;
    int64_t lo = 0, hi = 0;
    int rc = pml_peer_slice(PEER(peer), &lo, &hi);
    if (rc) { throw_for(env, rc); return; }
    jlong both[2] = { (jlong)lo, (jlong)hi };
    (*env)->SetLongArrayRegion(env, out, 0, 2, both);
}

NATIVE(jdouble, valueAndGradientAllReduce)(JNIEnv *env, jclass cls, jlong ds, jlong peer, jint loss, jdoubleArray coef,
                                           jdoubleArray outGrad, jboolean gatherOutput)
{
    (void)cls;
    double value = 0.0;
    double *w = pin(env, coef), *g = pin(env, outGrad);
    int rc = w ? pml_value_and_gradient_allreduce(DS(ds), PEER(peer), (int)loss, w, &value, g, gatherOutput ? 1 : 0) : PML_EINVAL;
    unpin_out(env, outGrad, g);
    unpin_in(env, coef, w);
    if (!w) { throw_msg(env, "java/lang/IllegalArgumentException", "null coefficients"); return 0.0; }
    if (rc) throw_for(env, rc);
    return (jdouble)value;
}

NATIVE(void, hessianVectorAllReduce)(JNIEnv *env, jclass cls, jlong ds, jlong peer, jint loss, jdoubleArray coef, jlong tag,
                                     jdoubleArray v, jdoubleArray out, jboolean gatherOutput)
{
    (void)cls;
    double *w = pin(env, coef), *x = pin(env, v), *h = pin(env, out);
    int rc = (x && h) ? pml_hessian_vector_allreduce(DS(ds), PEER(peer), (int)loss, w, (uint64_t)tag, x, h, gatherOutput ? 1 : 0)
                      : PML_EINVAL;
    unpin_out(env, out, h);
    unpin_in(env, v, x);
    unpin_in(env, coef, w);
    if (!(x && h)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

NATIVE(void, hessianDiagonalAllReduce)(JNIEnv *env, jclass cls, jlong ds, jlong peer, jint loss, jdoubleArray coef,
                                       jdoubleArray out, jboolean gatherOutput)
{
    (void)cls;
    double *w = pin(env, coef), *h = pin(env, out);
    int rc = (w && h) ? pml_hessian_diagonal_allreduce(DS(ds), PEER(peer), (int)loss, w, h, gatherOutput ? 1 : 0) : PML_EINVAL;
    unpin_out(env, out, h);
    unpin_in(env, coef, w);
    if (!(w && h)) { throw_msg(env, "java/lang/IllegalArgumentException", "null vector"); return; }
    if (rc) throw_for(env, rc);
}

NATIVE(void, peerDestroy)(JNIEnv *env, jclass cls, jlong peer)
{
    (void)cls;
    int rc = pml_peer_destroy(PEER(peer));
    if (rc) throw_for(env, rc);
}
