;;;; lowering.lisp — AbstractTensor view -> struct wf2cu_view.
;;;;
;;;; Replaces the loop nest `call-with-view` / `do-compiled-loop*` builds
;;;; (`source/vm/generic-tensor/call-with-view.lisp:24-229`, `do-compiled-loop.lisp:170-298`): instead
;;;; of handing a leaf ONE innermost row (`ViewInstruction{offset,size,by}`, `view.lisp:50-69`), the
;;;; whole view becomes kernel arguments:
;;;;     stride[k] = tensor-stride_k * (range-nth view_k 1  -  range-nth view_k 0)    ; 0 on :broadcast
;;;;     offset    = tensor-initial-offset + sum_k tensor-stride_k * (range-nth view_k 0)
;;;; exactly the arithmetic of `do-compiled-loop.lisp:203-214,219-236,254-295`, evaluated at call time
;;;; so adjustable symbols are already bound (`maybe-observe-axis`).
;;;; The Python twin is `CUDATensor.lower()` in cl-waffe2_b200/tensor.py, tested against the oracle's
;;;; restatement of the Lisp loop nest (tests/test_host_cpu.py).
(in-package :cl-waffe2/backends.cuda)

(declaim (inline observe))
(defun observe (x) (cl-waffe2/vm:maybe-observe-axis x))

(defun lower-view! (tensor view-ptr)
  "Fill the foreign wf2cu_view at VIEW-PTR from TENSOR.  Returns VIEW-PTR."
  (declare (type AbstractTensor tensor))
  (let* ((rank    (dims tensor))
	 (strides (tensor-stride tensor))
	 (views   (tensor-view tensor))
	 (offset  (tensor-initial-offset tensor)))
    (when (> rank +max-rank+)
      (error "CUDATensor: rank ~a exceeds WF2CU_MAX_RANK=~a" rank +max-rank+))
    (setf (foreign-slot-value view-ptr '(:struct wf2cu-view) 'base) (device-vec-ptr (tensor-vec tensor))
	  (foreign-slot-value view-ptr '(:struct wf2cu-view) 'rank) rank
	  (foreign-slot-value view-ptr '(:struct wf2cu-view) 'reserved) 0)
    (loop with shape-ptr  = (foreign-slot-pointer view-ptr '(:struct wf2cu-view) 'shape)
	  with stride-ptr = (foreign-slot-pointer view-ptr '(:struct wf2cu-view) 'stride)
	  for k fixnum from 0 below rank
	  for size   = (observe (nth k (shape tensor)))
	  for stride = (observe (nth k strides))
	  for sub    = (nth k views)
	  do (setf (mem-aref shape-ptr :int64 k) size)
	     (if (subscript-broadcast sub)
		 (setf (mem-aref stride-ptr :int64 k) 0)
		 (let* ((range (subscript-range sub))
			(first (wf/iter:range-nth range 0))
			(step  (- (wf/iter:range-nth range 1) first)))
		   (incf offset (* stride first))
		   (setf (mem-aref stride-ptr :int64 k) (* stride step)))))
    (setf (foreign-slot-value view-ptr '(:struct wf2cu-view) 'offset) offset)
    ;; A kernel trusts the descriptor: refuse one that addresses elements outside the storage (possible through
    ;; `permute*` of a sliced view, which re-applies the slice: `tensor.lisp:880-896,929-991`) instead of faulting.
    (let ((shape-ptr  (foreign-slot-pointer view-ptr '(:struct wf2cu-view) 'shape))
	  (stride-ptr (foreign-slot-pointer view-ptr '(:struct wf2cu-view) 'stride))
	  (lo offset)
	  (hi offset)
	  (empty-p nil))
      (dotimes (k rank)
	(let ((ext (mem-aref shape-ptr :int64 k))
	      (sd  (mem-aref stride-ptr :int64 k)))
	  (cond
	    ((<= ext 0) (setq empty-p t))
	    ((< sd 0)   (incf lo (* (1- ext) sd)))
	    (t          (incf hi (* (1- ext) sd))))))
      (when (and (not empty-p)
		 (or (< lo 0)
		     (>= (* 4 hi) (device-vec-bytes (tensor-vec tensor)))))
	(error "CUDATensor: a view of shape ~a addresses elements [~a, ~a] outside its storage of ~a elements"
	       (shape tensor) lo hi (floor (device-vec-bytes (tensor-vec tensor)) 4))))
    view-ptr))

(defmacro with-lowered-views ((&rest bindings) &body body)
  "(with-lowered-views ((xv x) (yv y)) ...) binds xv, yv to stack-allocated, filled wf2cu_view structs.
The counterpart of `with-tensor-ptrs` (`source/backends/cpu/tensor.lisp:43-57`)."
  (if (null bindings)
      `(progn ,@body)
      (destructuring-bind (var tensor) (car bindings)
	`(with-foreign-object (,var '(:struct wf2cu-view))
	   (tensor-vec ,tensor)		; make sure storage exists (lazy allocation)
	   (lower-view! ,tensor ,var)
	   (with-lowered-views ,(cdr bindings) ,@body)))))

(defun cuda-reject-p (&rest constructor-args)
  ":reject-p for every impl: decline when the device is unavailable or the node's dtype is not :float,
so dispatch falls to the next backend (`source/vm/nodes/defnode.lisp:215-232`)."
  (or (not *cuda-available-p*)
      (let ((dtype (find-if #'keywordp constructor-args)))
	(and dtype (not (eql dtype :float))))))

(defun cuda-only-reject-p (&rest constructor-args)
  ":reject-p for nodes whose constructor takes no dtype (`cuda-reject-p` would read a keyword argument as one)."
  (declare (ignore constructor-args))
  (not *cuda-available-p*))
