;;;; arithmetic.lisp — AddNode SubNode MulNode DivNode MoveTensorNode on CUDATensor.
;;;; Counterpart of `source/backends/cpu/arithmetic.lisp:65-142`.  `define-impl-op` (a plain lambda that
;;;; receives the live tensors each call, `source/vm/nodes/defnode.lisp:582-671`) is used instead of
;;;; `define-impl` because nothing needs compiling: the whole view is lowered at call time and ONE
;;;; kernel covers it (the CPU impl expands to a Lisp loop nest with one C call per innermost row).
;;;;
;;;; Reductions arrive here too: `!sum` is `A+=B out* x` where out* has (:broadcast n) axes
;;;; (`source/base-impl/reduction.lisp:60-79`); wf2cu_ew_binary classifies destination strides of 0 as
;;;; reduction axes and tree-reduces instead of racing.
(in-package :cl-waffe2/backends.cuda)

(macrolet ((define-arith-impl (node-name op-code)
	     `(define-impl-op (,node-name :device CUDATensor :reject-p #'cuda-reject-p)
		:forward ((self x y)
			  (declare (ignore self))
			  (with-lowered-views ((xv x) (yv y))
			    (ew-binary ,op-code xv yv))
			  x))))		; X <- X op Y, returns X as `vm.lisp:144-150` checks
  (define-arith-impl AddNode +add+)
  (define-arith-impl SubNode +sub+)
  (define-arith-impl MulNode +mul+)
  (define-arith-impl DivNode +div+))

(define-impl-op (MoveTensorNode :device CUDATensor :reject-p #'cuda-reject-p)
  :forward ((self x y)
	    ;; X <- Y unless the compiler pruned this copy (`base-impl/fundamental.lisp:9-43`)
	    (if (movetensor-ignore-me self)
		y
		(progn
		  (with-lowered-views ((xv x) (yv y))
		    (ew-binary +copy+ xv yv))
		  x))))

;; Scalar operands: `scalar (x) matrix` reaches the backend as an ordinary arithmetic node whose second operand is a
;; one-element tensor viewed with (:broadcast n) on every axis (`base-impl/arithmetic.lisp:28-51,244-250`): for a number
;; that tensor is made once at build time (`fill` in tensor.lisp), for a ScalarTensor it is refreshed by Scalar->MatNode's
;; `(setf vref)` (`fundamental.lisp:525-533`, a 4-byte transfer).  wf2cu_ew_binary hoists stride-0 operands into
;; registers, so nothing needs to be defined here; wf2cu_ew_scalar (scalar as a kernel argument) is bound in cffi.lisp
;; for hosts that know the number.
