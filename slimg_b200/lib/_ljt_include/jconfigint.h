#define BUILD "slimg-b200-handoff"
#undef inline
#define INLINE __inline__ __attribute__((always_inline))
#define THREAD_LOCAL __thread
#define PACKAGE_NAME "libjpeg-turbo"
#define VERSION "2.1.5.1"
#define SIZEOF_SIZE_T 8
#define HAVE_BUILTIN_CTZL
#define FALLTHROUGH __attribute__((fallthrough));
