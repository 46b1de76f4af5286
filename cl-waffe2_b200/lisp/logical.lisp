;;;; logical.lisp — Where-Operation-Node / Compare-Operation-Node on CUDATensor.
;;;; Counterpart of `source/backends/cpu/logical.lisp:37-84` (`waffe2_s{eq,lt,le,gt,ge}[_scal]`).
(in-package :cl-waffe2/backends.cuda)

(define-impl-op (Where-Operation-Node :device CUDATensor :reject-p #'cuda-reject-p)
  :forward ((self x out)
	    (let ((condition (car (logical-compiler-info self)))
		  (scal      (coerce (second (logical-compiler-info self)) 'single-float)))
	      (with-lowered-views ((xv x) (ov out))
		(cmp (cmp-code condition) xv (null-pointer) scal ov
		     (coerce (logical-true-then self) 'single-float)
		     (coerce (logical-false-then self) 'single-float)))
	      out)))

(define-impl-op (Compare-Operation-Node :device CUDATensor :reject-p #'cuda-reject-p)
  :forward ((self x y out)
	    (with-lowered-views ((xv x) (yv y) (ov out))
	      (cmp (cmp-code (logical-condition self)) xv yv 0.0 ov
		   (coerce (logical-true-then self) 'single-float)
		   (coerce (logical-false-then self) 'single-float)))
	    out))
