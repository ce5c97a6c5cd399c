/* NOT the JNI header.  A stand-in with the same names and shapes, only so that jni/oryx_b200_jni.c can be
 * syntax- and type-checked in an image that has no JDK (tests/test_jni_source.py: gcc -fsyntax-only).
 * A real build uses the JDK's <jni.h>. */
#ifndef ORYX_JNI_STUB_H_
#define ORYX_JNI_STUB_H_
#include <stdint.h>
#define JNIEXPORT
#define JNICALL
#define JNI_ABORT 2
#define JNI_TRUE 1
#define JNI_FALSE 0
typedef int32_t jint; typedef int64_t jlong; typedef int8_t jbyte; typedef uint8_t jboolean; typedef float jfloat;
typedef double jdouble; typedef jint jsize;
typedef void* jobject; typedef jobject jclass; typedef jobject jstring; typedef jobject jthrowable; typedef jobject jarray;
typedef jarray jintArray; typedef jarray jlongArray; typedef jarray jbyteArray; typedef jarray jfloatArray;
typedef jarray jdoubleArray; typedef jarray jobjectArray;
typedef struct _jmethodID* jmethodID;
struct JNINativeInterface_;
typedef const struct JNINativeInterface_* JNIEnv;
struct JNINativeInterface_ {
  jclass (*FindClass)(JNIEnv*, const char*);
  jint (*ThrowNew)(JNIEnv*, jclass, const char*);
  jint (*Throw)(JNIEnv*, jthrowable);
  jboolean (*ExceptionCheck)(JNIEnv*);
  jmethodID (*GetMethodID)(JNIEnv*, jclass, const char*, const char*);
  jobject (*NewObject)(JNIEnv*, jclass, jmethodID, ...);
  jstring (*NewStringUTF)(JNIEnv*, const char*);
  void (*DeleteLocalRef)(JNIEnv*, jobject);
  jsize (*GetArrayLength)(JNIEnv*, jarray);
  void* (*GetPrimitiveArrayCritical)(JNIEnv*, jarray, jboolean*);
  void (*ReleasePrimitiveArrayCritical)(JNIEnv*, jarray, void*, jint);
  jbyteArray (*NewByteArray)(JNIEnv*, jsize);
  void (*SetByteArrayRegion)(JNIEnv*, jbyteArray, jsize, jsize, const jbyte*);
  jintArray (*NewIntArray)(JNIEnv*, jsize);
  jlongArray (*NewLongArray)(JNIEnv*, jsize);
  jfloatArray (*NewFloatArray)(JNIEnv*, jsize);
  jdoubleArray (*NewDoubleArray)(JNIEnv*, jsize);
  jobjectArray (*NewObjectArray)(JNIEnv*, jsize, jclass, jobject);
  jobject (*GetObjectArrayElement)(JNIEnv*, jobjectArray, jsize);
  void (*SetObjectArrayElement)(JNIEnv*, jobjectArray, jsize, jobject);
  void (*SetIntArrayRegion)(JNIEnv*, jintArray, jsize, jsize, const jint*);
  void (*SetLongArrayRegion)(JNIEnv*, jlongArray, jsize, jsize, const jlong*);
  void (*SetFloatArrayRegion)(JNIEnv*, jfloatArray, jsize, jsize, const jfloat*);
  void (*SetDoubleArrayRegion)(JNIEnv*, jdoubleArray, jsize, jsize, const jdouble*);
  void (*GetFloatArrayRegion)(JNIEnv*, jfloatArray, jsize, jsize, jfloat*);
  void (*GetLongArrayRegion)(JNIEnv*, jlongArray, jsize, jsize, jlong*);
};
#endif
