#define JVERSION "6b  27-Mar-1998"
#define JCOPYRIGHT "see libjpeg-turbo LICENSE.md"
#define JCOPYRIGHT_SHORT "libjpeg-turbo"
