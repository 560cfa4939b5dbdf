/*
 * Minimal pthread-backed stand-in for the handful of SDL2 threading symbols that the
 * reference's Core/{ConcurrentDispatch,Debug,Thread,Semaphore}.cpp use.  TEST INFRASTRUCTURE
 * ONLY: lets the unmodified reference sources under /root/reference compile into
 * oracle/_ref/liboref_*.so without an SDL2 installation.  Not part of the product.
 */
#pragma once
#include <pthread.h>
#include <semaphore.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

typedef uint32_t Uint32;
typedef uint8_t Uint8;
typedef unsigned long SDL_threadID;

typedef struct SDL_mutex { pthread_mutex_t m; } SDL_mutex;
typedef struct SDL_cond { pthread_cond_t c; } SDL_cond;
typedef struct SDL_semaphore { sem_t s; } SDL_sem;
typedef int (*SDL_ThreadFunction)(void *);
typedef struct SDL_Thread { pthread_t t; SDL_ThreadFunction fn; void *data; int status; } SDL_Thread;

static inline SDL_mutex *SDL_CreateMutex(void) {
	SDL_mutex *m = (SDL_mutex *)malloc(sizeof(SDL_mutex));
	pthread_mutexattr_t a;
	pthread_mutexattr_init(&a);
	pthread_mutexattr_settype(&a, PTHREAD_MUTEX_RECURSIVE); /* SDL mutexes are recursive */
	pthread_mutex_init(&m->m, &a);
	pthread_mutexattr_destroy(&a);
	return m;
}
static inline void SDL_DestroyMutex(SDL_mutex *m) { if (m) { pthread_mutex_destroy(&m->m); free(m); } }
static inline int SDL_LockMutex(SDL_mutex *m) { return pthread_mutex_lock(&m->m); }
static inline int SDL_UnlockMutex(SDL_mutex *m) { return pthread_mutex_unlock(&m->m); }

static inline SDL_cond *SDL_CreateCond(void) {
	SDL_cond *c = (SDL_cond *)malloc(sizeof(SDL_cond));
	pthread_cond_init(&c->c, NULL);
	return c;
}
static inline void SDL_DestroyCond(SDL_cond *c) { if (c) { pthread_cond_destroy(&c->c); free(c); } }
static inline int SDL_CondWait(SDL_cond *c, SDL_mutex *m) { return pthread_cond_wait(&c->c, &m->m); }
static inline int SDL_CondSignal(SDL_cond *c) { return pthread_cond_signal(&c->c); }
static inline int SDL_CondBroadcast(SDL_cond *c) { return pthread_cond_broadcast(&c->c); }

static inline SDL_sem *SDL_CreateSemaphore(Uint32 v) {
	SDL_sem *s = (SDL_sem *)malloc(sizeof(SDL_sem));
	sem_init(&s->s, 0, v);
	return s;
}
static inline void SDL_DestroySemaphore(SDL_sem *s) { if (s) { sem_destroy(&s->s); free(s); } }
static inline int SDL_SemWait(SDL_sem *s) { return sem_wait(&s->s); }
static inline int SDL_SemPost(SDL_sem *s) { return sem_post(&s->s); }

static inline SDL_threadID SDL_ThreadID(void) { return (SDL_threadID)pthread_self(); }
static inline void *osb_sdl_thread_tramp(void *p) {
	SDL_Thread *t = (SDL_Thread *)p;
	t->status = t->fn(t->data);
	return NULL;
}
static inline SDL_Thread *SDL_CreateThread(SDL_ThreadFunction fn, const char *name, void *data) {
	(void)name;
	SDL_Thread *t = (SDL_Thread *)malloc(sizeof(SDL_Thread));
	t->fn = fn; t->data = data; t->status = 0;
	if (pthread_create(&t->t, NULL, osb_sdl_thread_tramp, t) != 0) { free(t); return NULL; }
	return t;
}
static inline void SDL_WaitThread(SDL_Thread *t, int *status) {
	if (!t) return;
	pthread_join(t->t, NULL);
	if (status) *status = t->status;
	free(t);
}

#define SDL_USEREVENT 0x8000
typedef struct SDL_UserEvent { Uint32 type; Uint32 timestamp; Uint32 windowID; int32_t code; void *data1; void *data2; } SDL_UserEvent;
typedef union SDL_Event { Uint32 type; SDL_UserEvent user; Uint8 padding[56]; } SDL_Event;
static inline int SDL_PushEvent(SDL_Event *e) { (void)e; return 1; }
