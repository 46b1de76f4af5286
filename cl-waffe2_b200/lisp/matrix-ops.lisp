;;;; matrix-ops.lisp — MatMulNode, MaxValue/MinValue/ArgMax/ArgMin on CUDATensor.
;;;; Counterpart of `source/backends/cpu/matrix-ops.lisp:12-369`.
(in-package :cl-waffe2/backends.cuda)

(defun batch-layout (tensor)
  "=> (values batch-count batch-stride) for the axes before the last two; they must collapse into one
stride (true after `adjust-layout`, which makes non-trivially viewed operands contiguous first)."
  (let* ((rank (dims tensor))
	 (nb   (- rank 2)))
    (if (<= nb 0)
	(values 1 0)
	(values (reduce #'* (mapcar #'observe (butlast (shape tensor) 2)))
		(if (subscript-broadcast (nth (1- nb) (tensor-view tensor)))
		    0
		    (observe (nth (1- nb) (tensor-stride tensor))))))))

(define-impl-op (MatMulNode :device CUDATensor :reject-p #'cuda-reject-p)
  :forward ((self a b out)
	    ;; C[~ i k] <- A[~ i j] . B[~ j k], alpha=1 beta=0.  `!t` operands arrive as flags; the
	    ;; untransposed storage is read through `read-untransposed` like the CPU impl does.
	    (let* ((ta (trans-a? self))
		   (tb (trans-b? self))
		   (a* (if ta (read-untransposed a) a))
		   (b* (if tb (read-untransposed b) b))
		   (m  (observe (car (last (shape out) 2))))
		   (n  (observe (car (last (shape out)))))
		   (k  (observe (car (last (shape a)))))) ; logical A is (~ m k)
	      (multiple-value-bind (batch sc) (batch-layout out)
		(multiple-value-bind (ba sa) (batch-layout a*)
		  (multiple-value-bind (bb sb) (batch-layout b*)
		    (with-lowered-views ((av a*) (bv b*) (cv out))
		      (flet ((off (v) (foreign-slot-value v '(:struct wf2cu-view) 'offset))
			     (ld  (v r) (mem-aref (foreign-slot-pointer v '(:struct wf2cu-view) 'stride) :int64 (- r 2))))
			(gemm (if ta 1 0) (if tb 1 0) m n k
			      (device-vec-ptr (tensor-vec a*)) (off av) (ld av (dims a*))
			      (device-vec-ptr (tensor-vec b*)) (off bv) (ld bv (dims b*))
			      (device-vec-ptr (tensor-vec out)) (off cv) (ld cv (dims out))
			      batch (if (= ba 1) 0 sa) (if (= bb 1) 0 sb) sc))))))
	      out)))

(macrolet ((define-rowred (node-name is-min arg-p)
	     `(define-impl-op (,node-name :device CUDATensor :reject-p #'cuda-reject-p)
		:forward ((self x out)
			  (declare (ignore self))
			  (with-lowered-views ((xv x) (ov out))
			    ,(if arg-p
				 `(argmaxmin ,is-min xv ov)
				 `(reduce-maxmin ,is-min xv ov)))
			  out))))
  (define-rowred MaxValue-Node 0 nil)
  (define-rowred MinValue-Node 1 nil)
  (define-rowred ArgMax-Node 0 t)
  (define-rowred ArgMin-Node 1 t))
