ssh test@127.0.0.1 "kill -9 3598"
ssh test@127.0.0.1 "kill -9 3599"
kill -9 2946
