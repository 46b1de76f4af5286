;;;; mathematics.lisp — unary math nodes and ExptNode on CUDATensor.
;;;; Counterpart of `source/backends/cpu/mathematics.lisp:8-75` (SLEEF u10 via `waffe2_s<f>`); the device
;;;; functions restate the same SLEEF polynomials (csrc/math.cuh), <= 2 ulp from the CPU path.
(in-package :cl-waffe2/backends.cuda)

(macrolet ((define-math-impl (node-name key)
	     `(define-impl-op (,node-name :device CUDATensor :reject-p #'cuda-reject-p)
		:forward ((self x out)
			  (declare (ignore self))
			  (with-lowered-views ((xv x) (ov out))
			    (ew-unary ,(getf *unary-codes* key) xv ov))
			  out))))
  (define-math-impl SinNode :sin)   (define-math-impl CosNode :cos)   (define-math-impl TanNode :tan)
  (define-math-impl ASinNode :asin) (define-math-impl ACosNode :acos) (define-math-impl ATanNode :atan)
  (define-math-impl SinHNode :sinh) (define-math-impl CosHNode :cosh) (define-math-impl TanHNode :tanh)
  (define-math-impl AbsNode :abs)   (define-math-impl ExpNode :exp)   (define-math-impl LogENode :loge)
  (define-math-impl Log2Node :log2) (define-math-impl Log10Node :log10) (define-math-impl Log1pNode :log1p)
  (define-math-impl SqrtNode :sqrt)
  ;; LispTensor-only in the reference (`base-impl/mathematics.lisp`), provided here so a model never
  ;; leaves the device for them:
  (define-math-impl SignNode :sign) (define-math-impl SquareNode :square))

(define-impl-op (ExptNode :device CUDATensor :reject-p #'cuda-reject-p)
  :forward ((self x out n)
	    (declare (ignore self))
	    ;; n is a ScalarTensor: a host number (`backends/cpu/mathematics.lisp:55-75`)
	    (with-lowered-views ((xv x) (ov out))
	      (pow xv ov (coerce (tensor-vec n) 'single-float)))
	    out))
