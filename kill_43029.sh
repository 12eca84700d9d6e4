ssh test@127.0.0.1 "kill -9 3582"
kill -9 2946
