;;;; package.lisp — cl-waffe2/backends.cuda
;;;;
;;;; The Common Lisp host side of the B200 `CUDATensor` backend.  A cl-waffe2 maintainer copies this
;;;; directory to `source/backends/cuda/` and adds the component list of INTEGRATION.md §2 to
;;;; `cl-waffe2.asd`.  Structure mirrors `source/backends/cpu/` file by file
;;;; (`package tensor arithmetic mathematics logical matrix-ops`), plus `cffi` (the binding that
;;;; `cl-waffe2-simd/api.lisp` is for the CPU backend) and `lowering` (view -> kernel arguments,
;;;; which replaces the `call-with-view` loop nest).
;;;;
;;;; NOTE: no Common Lisp implementation exists in the authoring image, so these files have been
;;;; written against the reference's source but never loaded.  The same host logic is exercised
;;;; through the identical C ABI by `cl-waffe2_b200/*.py`.
(in-package :cl-user)

(defpackage :cl-waffe2/backends.cuda
  (:documentation "B200 (sm_100a) device backend: CUDATensor -> CFFI -> libwaffe2cuda.so")
  (:use :cl :cl-waffe2/vm.generic-tensor :cl-waffe2/vm.nodes :cl-waffe2/base-impl :cffi)
  ;; cffi.lisp names each binding after its C entry point (wf2cu_fill -> `fill`, wf2cu_stream -> `stream`): these two
  ;; would otherwise redefine symbols of the locked COMMON-LISP package
  (:shadow #:fill #:stream)
  (:export
   #:CUDATensor
   #:*cuda-available-p*
   #:with-cuda
   #:cuda-sync
   #:cuda-comm-init
   #:cuda-allreduce-gradients
   #:cuda-layer-norm
   #:cuda-softmax
   #:cuda-self-attention))
(in-package :cl-waffe2/backends.cuda)

(defparameter *cuda-available-p* nil
  "T once libwaffe2cuda.so is loaded and wf2cu_init succeeded.  Every define-impl below has
`:reject-p` = (not *cuda-available-p*) or a non-:float dtype, so without a B200 dispatch falls through
to the next device in *using-backend* exactly like `*simd-extension-p*` does for CPUTensor
(`source/backends/cpu/package.lisp:48-78`).  There is no CPU code in this backend.")
